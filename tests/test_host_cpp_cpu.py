"""C++ host side above the C-ABI (include/usgscsm_b200_plugin.hpp, usgscsm_b200/host/cam_test.cpp):
it builds, parses options like the reference's tool (bin/usgscsm_cam_test.cc:47-119), resolves
model names without a device, and fails loudly -- no CPU fallback -- when there is no GPU."""
import ctypes as C
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOL = os.path.join(ROOT, "usgscsm_b200", "bin", "usgscsm_b200_cam_test")


def run(*args):
    return subprocess.run([TOOL] + list(args), capture_output=True, text=True, timeout=120)


def test_tool_is_built():
    assert os.access(TOOL, os.X_OK), "make -C usgscsm_b200 builds it"


def test_option_parsing_like_the_reference():
    r = run("--help")
    assert r.returncode == 1 and "Usage:" in r.stdout
    r = run("--model")
    assert r.returncode == 1 and "Could not parse correctly the input arguments." in r.stdout
    r = run("--sample-rate", "2")
    assert r.returncode == 1 and "The input model file is empty." in r.stdout
    r = run("--model", "x.json", "--sample-rate", "2.5")
    assert r.returncode == 1 and "The value of --sample-rate must be an integer." in r.stdout
    r = run("--model", "/nonexistent/x.json")
    assert r.returncode == 1 and "Cannot open file" in r.stdout


def test_no_device_fails_loudly(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from usgscsm_b200 import synth
    f = tmp_path / "state.json"
    f.write_text(synth.linescan_state(n_lines=64, n_samples=64))
    r = run("--model", str(f), "--sample-rate", "8")
    assert r.returncode == 1
    assert "Detected model state with model: USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL" in r.stdout
    assert "no usable CUDA device" in r.stderr
    assert "Norm of pixel errors" not in r.stdout


def test_model_name_from_state_needs_no_device():
    from usgscsm_b200 import _capi, synth
    buf = C.create_string_buffer(128)
    for state, name in ((synth.linescan_state(n_lines=64, n_samples=64), b"USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL"),
                        (synth.frame_state(), b"USGS_ASTRO_FRAME_SENSOR_MODEL"),
                        (synth.sar_state(), b"USGS_ASTRO_SAR_SENSOR_MODEL"),
                        (synth.pushframe_state(), b"USGS_ASTRO_PUSH_FRAME_SENSOR_MODEL")):
        raw = state.encode()
        assert _capi.lib.usgscsm_b200_model_name_from_state(raw, len(raw), buf, 128) == 0
        assert buf.value == name
    raw = b'{"m_modelName": "USGS_ASTRO_PROJECTED_SENSOR_MODEL"}'
    assert _capi.lib.usgscsm_b200_model_name_from_state(raw, len(raw), buf, 128) == _capi.EUNSUPPORTED
    raw = b'{"no_name": 1}'
    assert _capi.lib.usgscsm_b200_model_name_from_state(raw, len(raw), buf, 128) == _capi.ESTATE
    raw = b'USGS_ASTRO_FRAME_SENSOR_MODEL\n{"m_modelName": "USGS_ASTRO_FRAME_SENSOR_MODEL"}'
    assert _capi.lib.usgscsm_b200_model_name_from_state(raw, len(raw), buf, 8) == _capi.EINVAL
