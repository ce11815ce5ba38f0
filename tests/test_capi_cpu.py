"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/usgscsm_b200.h declares, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "usgscsm_b200.h")).read()
    return sorted(set(re.findall(r"USGSCSM_B200_API[^;]*?\b(usgscsm_b200_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from usgscsm_b200 import _capi
    names = _declared()
    assert len(names) >= 19
    lib = ctypes.CDLL(_capi.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), n
    assert sorted(_capi.SIGNATURES) == names
    assert _capi.lib.usgscsm_b200_abi_version() == 1


def test_status_bits_match_header():
    from usgscsm_b200 import _capi
    text = open(os.path.join(ROOT, "include", "usgscsm_b200.h")).read()
    for name, shift in re.findall(r"#define USGSCSM_B200_(ST_\w+) \(1u << (\d+)\)", text):
        assert getattr(_capi, name) == 1 << int(shift)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, synth
    assert _capi.lib.usgscsm_b200_device_count() == 0
    state = synth.linescan_state(n_lines=64, n_samples=64)
    with pytest.raises(_capi.EngineError) as e:
        ub.UsgsAstroPlugin().constructModelFromState(state)
    assert e.value.code == _capi.ENODEVICE


def test_product_does_not_touch_the_oracle():
    """Nothing under usgscsm_b200/ or include/ may name the oracle."""
    for base in ("usgscsm_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, base)):
            if "build" in dirpath or "__pycache__" in dirpath:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                    src = open(os.path.join(dirpath, f), errors="replace").read()
                    assert "oracle" not in src.lower() or f == "_capi.py", os.path.join(dirpath, f)


def test_plugin_metadata_and_model_name():
    import usgscsm_b200 as ub
    from usgscsm_b200 import synth
    p = ub.UsgsAstroPlugin()
    assert p.getPluginName() == "UsgsAstroPluginCSM"
    assert p.getManufacturer() == "UsgsAstrogeology"
    assert p.getReleaseDate() == "20190222"
    state = synth.linescan_state(n_lines=64, n_samples=64)
    assert p.getModelNameFromModelState(state) == "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL"
    assert p.canModelBeConstructedFromState("USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL", state)
    assert not p.canModelBeConstructedFromState("USGS_ASTRO_FRAME_SENSOR_MODEL", state)
    with pytest.raises(ub.Error):
        p.getModelNameFromModelState("{\"no_name\": 1}")
