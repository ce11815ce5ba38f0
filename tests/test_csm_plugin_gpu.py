"""libusgscsm_b200_csm.so as a REAL CSM plugin: a program that knows only the
CSM API (usgscsm_b200/host/csm_plugin_test.cpp; links libcsmapi, includes csm/*.h) dlopen()s the
library, finds the plugin with csm::Plugin::findPlugin("UsgsAstroPluginCSM"), builds models from
states and from ISDs and calls the csm::RasterGM virtuals one point at a time.  Its per-point
results are compared with the reference's (oracle) on the same state, and the state the plugin's
model returns must load in the REFERENCE plugin and give the same answers."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

import ref_goldens as G

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "usgscsm_b200", "libusgscsm_b200_csm.so")
TOOL = os.path.join(ROOT, "usgscsm_b200", "bin", "csm_plugin_test")


def _run(kind, path, pts):
    if not (os.path.exists(PLUGIN) and os.path.exists(TOOL)):
        pytest.fail("libusgscsm_b200_csm.so / csm_plugin_test not built (make -C usgscsm_b200)")
    with tempfile.TemporaryDirectory() as d:
        pin, pout = os.path.join(d, "pts.bin"), os.path.join(d, "out.bin")
        np.ascontiguousarray(pts, dtype=np.float64).tofile(pin)
        r = subprocess.run([TOOL, PLUGIN, kind, path, pin, str(len(pts)), pout],
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout + r.stderr
        raw = open(pout, "rb").read()
    n = len(pts)
    out = np.frombuffer(raw[:n * 80], dtype=np.float64).reshape(n, 10)
    state = raw[n * 80:].decode("utf-8")
    return out, state, r.stdout


def _points(rng, nl, ns, n):
    return np.stack([rng.uniform(0, nl, n), rng.uniform(0, ns, n), rng.uniform(-500, 500, n)], axis=1)


def _compare(ref, out, state, pts, px_tol=1e-6):
    rm = ref.RefModel(state)   # the plugin's getModelState() loads in the REFERENCE plugin
    x, y, z, a, st = rm.image_to_ground(pts[:, 0], pts[:, 1], pts[:, 2])
    err = out[:, 9] != 0
    assert np.array_equal(err, (st & 0x260) != 0)            # thrown csm::Error <-> error bits
    ok = ~err
    assert max(np.abs(out[ok, 0] - x[ok]).max(), np.abs(out[ok, 1] - y[ok]).max(),
               np.abs(out[ok, 2] - z[ok]).max()) <= 1e-4
    fin = ok & np.isfinite(a)
    assert np.abs(out[fin, 3] - a[fin]).max(initial=0) <= 1e-4
    assert np.array_equal(out[ok, 4].astype(int), (st[ok] & 3))          # warnings
    l, s, a2, st2 = rm.ground_to_image(out[ok, 0], out[ok, 1], out[ok, 2])
    assert max(np.abs(out[ok, 5] - l).max(), np.abs(out[ok, 6] - s).max()) <= px_tol
    assert np.array_equal(out[ok, 8].astype(int), (st2 & 3))


def test_findplugin_state_sup(ref):
    """The reference-held HiRISE .sup state, through dlopen -> findPlugin ->
    constructModelFromState -> imageToGround / groundToImage per point."""
    rng = np.random.default_rng(3)
    pts = _points(rng, 40000, 20000, 400)
    out, state, log = _run("state", os.path.join(G.DATA, "gxp_model_file.sup"), pts)
    assert "plugin UsgsAstroPluginCSM by UsgsAstrogeology, 4 models" in log
    assert "model USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL family GeometricRaster pedigree USGS_LINE_SCANNER" in log
    assert "image 40000 x 20000 params 16 [IT Pos. Bias   ] sensor EO/PUSHBROOM" in log
    _compare(ref, out, state, pts)


@pytest.mark.parametrize("kind", ["pushframe", "frame", "sar"])
def test_findplugin_state_other_models(ref, kind):
    from usgscsm_b200 import synth
    state = {"pushframe": lambda: synth.pushframe_state(n_framelets=64, framelet_height=16, n_samples=1024),
             "frame": lambda: synth.frame_state(n_lines=2048, n_samples=2048),
             "sar": lambda: synth.sar_state()}[kind]()
    rm = ref.RefModel(state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(4)
    pts = _points(rng, nl, ns, 300)
    with tempfile.NamedTemporaryFile("w", suffix=".json", delete=False) as f:
        f.write(state)
    try:
        out, st, log = _run("state", f.name, pts)
    finally:
        os.unlink(f.name)
    assert {"pushframe": "USGS_ASTRO_PUSH_FRAME", "frame": "USGS_ASTRO_FRAME", "sar": "USGS_ASTRO_SAR"}[kind] in log
    _compare(ref, out, st, pts, px_tol=1e-6)


@pytest.mark.parametrize("name", ["constVelocityLineScan", "orbitalLineScan", "orbitalPushFrame",
                                  "orbitalFramer", "orbitalSar", "toughLroNacLineScan"])
def test_findplugin_isd(ref, name):
    """constructModelFromISD through the CSM API, on the reference's ISD fixtures; the reference
    date of the synthetic fixtures is what the reference's tests assert
    (LineScanCameraTests.cpp:274-277, SarTests.cpp:257-260)."""
    rm = ref.RefModel.from_isd(G.isd_image(name))
    nl, ns = rm.image_size()
    rng = np.random.default_rng(5)
    pts = _points(rng, nl, ns, 200)
    pts[:, 2] = rng.uniform(-1, 1, len(pts)) * (1.0 if "constVelocity" in name else 100.0)
    out, state, log = _run("isd", G.isd_image(name), pts)
    assert "id %s " % name in log
    if name == "orbitalLineScan":
        assert "date 2000-01-01T12:15:34.816Z" in log
    if name == "orbitalSar":
        assert "date 2020-08-16T20:51:08.816Z" in log
    _compare(ref, out, state, pts, px_tol=10.0 if "PushFrame" in name else 1e-6)
