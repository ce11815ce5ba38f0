"""Tuning knobs are read once per process, so each variant runs in its own interpreter: every
combination must compute the same bits."""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(code, **env):
    e = dict(os.environ, **{k: str(v) for k, v in env.items()})
    r = subprocess.run([sys.executable, "-c", "import sys; sys.path.insert(0, %r)\n" % ROOT + textwrap.dedent(code)],
                       env=e, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r.stdout


def test_grid_chunks_compute_the_same_lines():
    """a dense grid cut into pipeline chunks must give every row the line
    line0 + (row0 + r) * dline, whatever the chunk size -- non-dyadic dline, 4 Ki-point chunks."""
    out = _run("""
        import numpy as np, torch
        import usgscsm_b200 as ub
        from usgscsm_b200 import synth
        gm = ub.UsgsAstroPlugin().constructModelFromState(synth.linescan_state(n_lines=4096, n_samples=512, seed=3))
        nl, ns = 700, 37
        h = np.linspace(-300.0, 300.0, nl * ns)
        x, y, z, a, st = gm.imageToGroundGrid(0.3, 0.7, 0.1, 1.3, nl, ns, h, achieved=True)      # host, chunked
        dx, dy, dz, da, dst = gm.imageToGroundGrid(0.3, 0.7, 0.1, 1.3, nl, ns, torch.as_tensor(h, device="cuda:0"),
                                                   achieved=True)                                 # device, one launch
        line = np.repeat(0.3 + np.arange(nl) * 0.1, ns); samp = np.tile(0.7 + np.arange(ns) * 1.3, nl)
        ex, ey, ez, ea, est = gm.imageToGroundBatch(line, samp, h)                                # explicit pixels
        for u, v, w in ((x, dx, ex), (y, dy, ey), (z, dz, ez), (a, da, ea)):
            assert np.array_equal(u, v.cpu().numpy()) and np.array_equal(u, w)
        assert np.array_equal(st, dst.cpu().numpy().view(np.uint32)) and np.array_equal(st, est)
        print("OK")
        """, USGSCSM_B200_PIPELINE_CHUNK=4096)
    assert "OK" in out


@pytest.mark.parametrize("env", [dict(USGSCSM_B200_SAR_REFILL=1), dict(USGSCSM_B200_SAR_REFILL=1, USGSCSM_B200_SAR_PAD=1),
                                 dict(USGSCSM_B200_SAR_PAD=1, USGSCSM_B200_SAR_MINB=2),
                                 dict(USGSCSM_B200_SAR_INTERLEAVE=0), dict(USGSCSM_B200_SAR_INTERLEAVE=0, USGSCSM_B200_SAR_MINB=2)])
def test_sar_kernel_variants_same_bits(env):
    """Lane-refill / padded-table variants of SAR groundToImage against the
    default kernel's output saved by the first run."""
    code = """
        import numpy as np, torch, os
        import usgscsm_b200 as ub
        from usgscsm_b200 import synth
        gm = ub.UsgsAstroPlugin().constructModelFromState(synth.sar_state())
        rng = np.random.default_rng(9)
        n = 200000
        line, samp, h = rng.uniform(0, 64576, n), rng.uniform(0, 2367, n), rng.uniform(-900, 900, n)
        x, y, z, _, _ = gm.imageToGroundBatch(line, samp, h)
        x[:500] *= 1.3    # no Doppler bracket / far off: the error paths
        l, s, a, st = gm.groundToImageBatch(x, y, z)
        np.save(os.environ["OUT"], np.concatenate([l, s, st.astype(np.float64)]))
        print("OK")
        """
    import tempfile
    import numpy as np
    with tempfile.TemporaryDirectory() as d:
        _run(code, OUT=os.path.join(d, "a.npy"))
        _run(code, OUT=os.path.join(d, "b.npy"), **env)
        a, b = np.load(os.path.join(d, "a.npy")), np.load(os.path.join(d, "b.npy"))
        # the evaluation counter (status bits 16-23) is the same too: same search, same order
        assert np.array_equal(a, b, equal_nan=True)


def test_pageable_and_pinned_arrays_same_bits():
    """numpy (pageable) arrays go through the engine's page-locked ring with copier threads;
    pinned arrays through the three-slot pipeline: same results, several chunks each."""
    out = _run("""
        import numpy as np, torch
        import usgscsm_b200 as ub
        from usgscsm_b200 import synth, _capi
        gm = ub.UsgsAstroPlugin().constructModelFromState(synth.linescan_state(n_lines=4096, n_samples=1024, seed=8))
        n = 300000
        rng = np.random.default_rng(2)
        line, samp, h = rng.uniform(0, 4096, n), rng.uniform(0, 1024, n), rng.uniform(-500, 500, n)
        x, y, z, a, st = gm.imageToGroundBatch(line, samp, h)                 # pageable in / out
        pin = lambda v: torch.as_tensor(v).pin_memory()
        pl, ps, ph = pin(line), pin(samp), pin(h)
        px, py, pz, pa = (torch.empty(n, dtype=torch.float64).pin_memory() for _ in range(4))
        pst = torch.empty(n, dtype=torch.int32).pin_memory()
        _capi.check(_capi.lib.usgscsm_b200_image_to_ground(gm._h, n, pl.data_ptr(), ps.data_ptr(), ph.data_ptr(), 1, 0.001,
                                                           px.data_ptr(), py.data_ptr(), pz.data_ptr(), pa.data_ptr(), pst.data_ptr()))
        assert np.array_equal(x, px.numpy()) and np.array_equal(z, pz.numpy()) and np.array_equal(a, pa.numpy())
        assert np.array_equal(st, pst.numpy().view(np.uint32))
        l, s, ach, st2 = gm.groundToImageBatch(x, y, z)
        assert np.abs(l - line).max() < 5e-3 and np.abs(s - samp).max() < 5e-3
        print("OK")
        """, USGSCSM_B200_STAGE_CHUNK=20000, USGSCSM_B200_STAGE_THREADS=5, USGSCSM_B200_PIPELINE_CHUNK=50000)
    assert "OK" in out


def test_fused_all_sensor_partials_same_bits():
    """USGSCSM_B200_LS_FUSED_PARTIALS=1: computeAllSensorPartials in one kernel (shared seed and
    first two evaluations) against the default per-parameter searches, unadjusted and adjusted
    model, both overloads."""
    code = """
        import numpy as np, os
        import usgscsm_b200 as ub
        from usgscsm_b200 import synth
        gm = ub.UsgsAstroPlugin().constructModelFromState(synth.linescan_state(n_lines=4096, n_samples=1024, seed=12))
        rng = np.random.default_rng(3)
        n = 20000
        line, samp, h = rng.uniform(0, 4096, n), rng.uniform(0, 1024, n), rng.uniform(-500, 500, n)
        x, y, z, _, _ = gm.imageToGroundBatch(line, samp, h)
        out = []
        for params in ({}, {1: 1.5, 5: 0.01, 8: -3e-5, 15: 0.01}):
            for i, v in params.items():
                gm.setParameterValue(i, v)
            p, a, st = gm.computeAllSensorPartialsBatch(x, y, z)
            p2, a2, st2 = gm.computeAllSensorPartialsBatch(x, y, z, imagePts=(line, samp), indices=[0, 7, 15])
            out += [p.reshape(-1), a, st.astype(np.float64), p2.reshape(-1), a2, st2.astype(np.float64)]
        np.save(os.environ["OUT"], np.concatenate(out))
        print("OK")
        """
    import tempfile
    import numpy as np
    with tempfile.TemporaryDirectory() as d:
        _run(code, OUT=os.path.join(d, "a.npy"))
        _run(code, OUT=os.path.join(d, "b.npy"), USGSCSM_B200_LS_FUSED_PARTIALS=1)
        a, b = np.load(os.path.join(d, "a.npy")), np.load(os.path.join(d, "b.npy"))
        assert np.array_equal(a, b, equal_nan=True)


def test_frame_kernel_variants_same_bits():
    """The frame kernels specialised on the distortion type, two points per 128-bit access
    (default) against the generic one-point kernel; odd lengths, arrays off the 16-byte
    boundary, radial / transverse / other distortion, sphere and oblate body, dense grid."""
    code = """
        import numpy as np, os, torch
        import usgscsm_b200 as ub
        from usgscsm_b200 import synth
        out = []
        dev = torch.device("cuda:0")
        for kw in (dict(), dict(distortion_type=synth.TRANSVERSE, dist_coeffs=(0.0, 1.0, 0.0, 1e-6, 2e-6, -1e-6, 1e-8, 0.0, 2e-8, 0.0,
                                                                       0.0, 0.0, 1.0, -1e-6, 1e-6, 2e-6, 0.0, 1e-8, 0.0, 2e-8)),
                   dict(distortion_type=synth.DAWNFC, dist_coeffs=(8.4e-6,)), dict(minor_axis=synth.MOON_RADIUS * 0.99),
                   dict(summing=2, starting_sample=3.0)):
            gm = ub.UsgsAstroPlugin().constructModelFromState(synth.frame_state(n_lines=1024, n_samples=766, **kw))
            nl, ns = 301, 766
            n = nl * ns
            h = torch.linspace(-200.0, 300.0, n + 3, dtype=torch.float64, device=dev)
            g = gm.imageToGroundGrid(0.5, 0.5, 1.0, 1.0, nl, ns, h[:n], achieved=True)
            out += [t.double().cpu().numpy() for t in g]
            for off, m in ((0, n), (1, n - 1), (0, n - 1), (1, n - 2)):   # aligned / shifted by 8 bytes, even / odd
                x, y, z = (t[off:off + m] for t in g[:3])
                r = gm.groundToImageBatch(x, y, z, achieved=True)
                out += [t.double().cpu().numpy() for t in r]
                b = gm.imageToGroundBatch(r[0], r[1], h[off:off + m], achieved=True)
                out += [t.double().cpu().numpy() for t in b]
        np.save(os.environ["OUT"], np.concatenate(out))
        print("OK")
        """
    import tempfile
    import numpy as np
    with tempfile.TemporaryDirectory() as d:
        _run(code, OUT=os.path.join(d, "a.npy"))
        _run(code, OUT=os.path.join(d, "b.npy"), USGSCSM_B200_FR_PAIRS=0)
        _run(code, OUT=os.path.join(d, "c.npy"), USGSCSM_B200_FR_SPECIALISE=0)
        a, b, c = (np.load(os.path.join(d, f)) for f in ("a.npy", "b.npy", "c.npy"))
        assert a.size > 10 ** 6
        assert np.array_equal(a, b, equal_nan=True) and np.array_equal(a, c, equal_nan=True)


def test_pushframe_kernel_variants_same_bits():
    """The push-frame kernels specialised on distortion type / attitude correction (framelet
    numbers without a division, hoisted line split, orientation outside the table out of line)
    against the generic kernels (USGSCSM_B200_PF_SPECIALISE=0): lines far outside the image,
    negative lines, flipped / reversed framelets, summing, oblate body, adjusted parameters."""
    code = """
        import numpy as np, os, torch
        import usgscsm_b200 as ub
        from usgscsm_b200 import synth
        T = (0.0, 1.0, 0.0, 1e-6, 2e-6, -1e-6, 1e-8, 0.0, 2e-8, 0.0, 0.0, 0.0, 1.0, -1e-6, 1e-6, 2e-6, 0.0, 1e-8, 0.0, -2e-8)
        out = []
        rng = np.random.default_rng(8)
        for kw in (dict(), dict(flipped=True, reversed_order=True, line_sign=1.0), dict(summing=2.0),
                   dict(distortion_type=1, dist_coeffs=T), dict(distortion_type=3, dist_coeffs=(8.4e-6,)),
                   dict(major_axis=3396190.0, minor_axis=3376200.0, gm=4.2828e13, altitude=300000.0),
                   dict(params=(3.0, -2.0, 1.5, 0.02, -0.01, 0.005, 0.5, -0.3, 0.2, 0.05, 0.04, -0.03, 1e-4, 2e-4, -1e-4, 0.01)),
                   dict(n_framelets=3, framelet_height=7, n_samples=200)):
            gm = ub.UsgsAstroPlugin().constructModelFromState(synth.pushframe_state(n_framelets=kw.pop("n_framelets", 64),
                                                               framelet_height=kw.pop("framelet_height", 16),
                                                               n_samples=kw.pop("n_samples", 600), seed=4, **kw))
            sz = gm.getImageSize(); nl, ns = int(sz.line), int(sz.samp)
            g = gm.imageToGroundGrid(0.5, 0.5, 1.0, 1.0, nl, ns, np.linspace(-300.0, 300.0, nl * ns), achieved=True)
            out += [np.asarray(t, dtype=np.float64) for t in g]
            n = 40000
            line = np.concatenate([rng.uniform(-3 * nl, 4 * nl, n), np.arange(-40.0, 40.0, 0.25), nl + np.arange(-40.0, 40.0, 0.25),
                                   [1e300, -1e300, 5e15, -5e15, 4503599627370496.0, np.nan, np.inf, -0.0, 0.0]])
            samp = rng.uniform(-50, ns + 50, line.size)
            b = gm.imageToGroundBatch(line, samp, rng.uniform(-500, 500, line.size), achieved=True)
            out += [np.asarray(t, dtype=np.float64) for t in b]
            r = gm.groundToImageBatch(g[0], g[1], g[2], achieved=True)
            out += [np.asarray(t, dtype=np.float64) for t in r]
        np.save(os.environ["OUT"], np.concatenate(out))
        print("OK")
        """
    import tempfile
    import numpy as np
    with tempfile.TemporaryDirectory() as d:
        _run(code, OUT=os.path.join(d, "a.npy"))
        _run(code, OUT=os.path.join(d, "b.npy"), USGSCSM_B200_PF_SPECIALISE=0)
        a, b = np.load(os.path.join(d, "a.npy")), np.load(os.path.join(d, "b.npy"))
        assert a.size > 10 ** 6
        assert np.array_equal(a, b, equal_nan=True)
