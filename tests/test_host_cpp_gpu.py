"""The C++ host mirror end to end on the GPU: usgscsm_b200_cam_test (the reference's
bin/usgscsm_cam_test.cc loop as two batch calls through include/usgscsm_b200_plugin.hpp and the
AoS C-ABI entry points) against the reference's per-point calls on the same pixels, and the
AoS entry points against the SoA ones."""
import os
import subprocess

import numpy as np
import pytest

from parity_util import M_TOL, PX_TOL

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOOL = os.path.join(ROOT, "usgscsm_b200", "bin", "usgscsm_b200_cam_test")


def _states():
    from usgscsm_b200 import synth
    return {
        "linescan": synth.linescan_state(n_lines=512, n_samples=256, seed=5),
        "pushframe": synth.pushframe_state(n_framelets=16, framelet_height=16, n_samples=256),
        "frame": synth.frame_state(n_lines=256, n_samples=256),
        "sar": synth.sar_state(),
    }


@pytest.mark.parametrize("kind", ["linescan", "pushframe", "frame", "sar"])
def test_cam_test_tool_matches_reference_per_point_calls(ref, tmp_path, kind):
    state = _states()[kind]
    f = tmp_path / "state.json"
    f.write_text(state)
    pts = tmp_path / "points.txt"
    rate = 64 if kind == "sar" else 8
    r = subprocess.run([TOOL, "--model", str(f), "--sample-rate", str(rate), "--subpixel-offset", "0.25",
                        "--height-above-datum", "100", "--desired-precision", "1e-8",
                        "--points-file", str(pts)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Norm of pixel errors after projecting from camera to ground and back." in r.stdout
    rows = [[float.fromhex(t) if i < 9 else int(t) for i, t in enumerate(l.split())]
            for l in pts.read_text().splitlines()]
    a = np.array(rows, dtype=np.float64)
    rm = ref.RefModel(state)
    if kind == "linescan":
        import usgscsm_b200 as ub
        gm = ub.UsgsAstroPlugin().constructModelFromState(state)
        rm.ls_set_projective(True, gm.lsDerived()[5:19])
    nl, ns = rm.image_size()
    # the reference's loop order: samples outer, lines inner (bin/usgscsm_cam_test.cc:294-296)
    exp = [(l + 0.25, s + 0.25) for s in range(0, ns, rate) for l in range(0, nl, rate)]
    assert len(exp) == len(a) and ("Count:  %d" % len(exp)) in r.stdout
    assert np.array_equal(a[:, 0], [e[0] for e in exp]) and np.array_equal(a[:, 1], [e[1] for e in exp])
    rx, ry, rz, ra, rst = rm.image_to_ground(a[:, 0], a[:, 1], np.full(len(a), 100.0), 1e-8)
    assert max(np.abs(rx - a[:, 2]).max(), np.abs(ry - a[:, 3]).max(), np.abs(rz - a[:, 4]).max()) <= M_TOL
    fa = np.isfinite(ra)   # models that never write achievedPrecision report NaN on both sides
    assert np.array_equal(np.isfinite(a[:, 5]), fa)
    assert np.abs(ra[fa] - a[fa, 5]).max(initial=0.0) <= M_TOL
    assert np.array_equal(rst, a[:, 9].astype(np.uint32))
    rl, rs, ra2, rst2 = rm.ground_to_image(a[:, 2].copy(), a[:, 3].copy(), a[:, 4].copy(), 1e-8)
    assert max(np.abs(rl - a[:, 6]).max(), np.abs(rs - a[:, 7]).max()) <= PX_TOL
    assert np.array_equal(rst2, a[:, 10].astype(np.uint32))
    err = np.sort(np.hypot(a[:, 0] - a[:, 6], a[:, 1] - a[:, 7]))
    assert ("Max:    %s" % _cxx_double(err[-1])) in r.stdout


def _cxx_double(v):
    """operator<< of a double at the default precision (6 significant digits, %g)."""
    return "%g" % v


@pytest.mark.parametrize("kind", ["linescan", "frame"])
def test_aos_entry_points_equal_soa(kind):
    import ctypes as C
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi
    gm = ub.UsgsAstroPlugin().constructModelFromState(_states()[kind])
    sz = gm.getImageSize()
    rng = np.random.default_rng(3)
    n = 5_000_000 if kind == "frame" else 100_003   # frame: more than one pipeline chunk
    ls = np.stack([rng.uniform(0, sz.line, n), rng.uniform(0, sz.samp, n)], axis=1).copy()
    h = rng.uniform(-500, 500, n)
    xyz = np.empty((n, 3))
    ach = np.empty(n)
    st = np.empty(n, dtype=np.uint32)
    _capi.check(_capi.lib.usgscsm_b200_image_to_ground_aos(
        gm._h, n, ls.ctypes.data, h.ctypes.data, 1, 0.001, xyz.ctypes.data, ach.ctypes.data, st.ctypes.data))
    x, y, z, a, s = gm.imageToGroundBatch(ls[:, 0].copy(), ls[:, 1].copy(), h)
    assert np.array_equal(xyz[:, 0], x) and np.array_equal(xyz[:, 1], y) and np.array_equal(xyz[:, 2], z)
    assert np.array_equal(ach, a, equal_nan=True) and np.array_equal(st, s)
    back = np.empty((n, 2))
    _capi.check(_capi.lib.usgscsm_b200_ground_to_image_aos(
        gm._h, n, xyz.ctypes.data, 0.001, back.ctypes.data, ach.ctypes.data, st.ctypes.data))
    l, sm, a2, s2 = gm.groundToImageBatch(x, y, z)
    assert np.array_equal(back[:, 0], l) and np.array_equal(back[:, 1], sm)
    assert np.array_equal(ach, a2, equal_nan=True) and np.array_equal(st, s2)
    # one height for all points, no optional outputs
    h1 = np.array([25.0])
    _capi.check(_capi.lib.usgscsm_b200_image_to_ground_aos(
        gm._h, 1000, ls.ctypes.data, h1.ctypes.data, 0, 0.001, xyz.ctypes.data, None, None))
    x, y, z, _, _ = gm.imageToGroundBatch(ls[:1000, 0].copy(), ls[:1000, 1].copy(), 25.0)
    assert np.array_equal(xyz[:1000, 0], x) and np.array_equal(xyz[:1000, 2], z)
    assert _capi.lib.usgscsm_b200_ground_to_image_aos(gm._h, 0, None, 0.001, None, None, None) == 0


def test_cam_test_binary_state_round_trip(ref, tmp_path):
    """--output-binary-model-state writes the reference's msgpack bytes; the tool then loads
    that file and reports the same errors as from the JSON state."""
    from oracle import ref as oracle_ref
    state = _states()["linescan"]
    f = tmp_path / "state.json"
    f.write_text(state)
    packed = tmp_path / "state.bin"
    args = ["--sample-rate", "32", "--desired-precision", "1e-8"]
    r1 = subprocess.run([TOOL, "--model", str(f), "--output-binary-model-state", str(packed)] + args,
                        capture_output=True, text=True, timeout=300)
    assert r1.returncode == 0, r1.stdout + r1.stderr
    # the tool packs getModelState() -- the state with the values updateState derived at load
    # written back, as the reference's tool does; the bytes are nlohmann's for that document
    import usgscsm_b200 as ub
    current = ub.UsgsAstroPlugin().constructModelFromState(state).getModelState()
    assert packed.read_bytes() == oracle_ref.state_to_msgpack(current)
    r2 = subprocess.run([TOOL, "--model", str(packed)] + args, capture_output=True, text=True, timeout=300)
    assert r2.returncode == 0, r2.stdout + r2.stderr
    assert "Detected msgpack binary model state" in r2.stdout
    pick = lambda out: [l for l in out.splitlines() if l.startswith(("Min:", "Median:", "Max:", "Count:"))]
    assert pick(r1.stdout) == pick(r2.stdout) and len(pick(r1.stdout)) == 4
