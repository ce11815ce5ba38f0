"""usgscsm_b200_*_multi (one process, several models / devices, contiguous chunks, no exchange
step) and concurrent calls from several host threads on different models (SURVEY.md §8b
"Threading").  With one GPU the models all live on device 0; with more they are spread."""
import ctypes as C
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _models(state, count):
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi
    ndev = _capi.lib.usgscsm_b200_device_count()
    return [ub.UsgsAstroPlugin(device=g % ndev).constructModelFromState(state) for g in range(count)]


@pytest.mark.parametrize("count", [1, 2, 3])
def test_multi_equals_single(count):
    from usgscsm_b200 import _capi, synth
    state = synth.linescan_state(n_lines=4096, n_samples=512, seed=21)
    gms = _models(state, count)
    handles = (C.c_void_p * count)(*[g._h for g in gms])
    rng = np.random.default_rng(1)
    n = 1_000_003
    line, samp = rng.uniform(0, 4096, n), rng.uniform(0, 512, n)
    h = synth.synthetic_dem(line, samp)
    x, y, z, a = (np.empty(n) for _ in range(4))
    st = np.empty(n, dtype=np.uint32)
    _capi.check(_capi.lib.usgscsm_b200_image_to_ground_multi(
        handles, count, n, line.ctypes.data, samp.ctypes.data, h.ctypes.data, 1, 0.001,
        x.ctypes.data, y.ctypes.data, z.ctypes.data, a.ctypes.data, st.ctypes.data))
    ex, ey, ez, ea, est = gms[0].imageToGroundBatch(line, samp, h)
    assert np.array_equal(x, ex) and np.array_equal(y, ey) and np.array_equal(z, ez)
    assert np.array_equal(a, ea) and np.array_equal(st, est)
    l, s, a2 = (np.empty(n) for _ in range(3))
    _capi.check(_capi.lib.usgscsm_b200_ground_to_image_multi(
        handles, count, n, x.ctypes.data, y.ctypes.data, z.ctypes.data, 0.001,
        l.ctypes.data, s.ctypes.data, a2.ctypes.data, st.ctypes.data))
    el, es, ea2, est2 = gms[0].groundToImageBatch(x, y, z)
    assert np.array_equal(l, el) and np.array_equal(s, es)
    assert np.array_equal(a2, ea2) and np.array_equal(st, est2)
    # one height for every point, optional outputs absent, fewer points than models
    h1 = np.array([10.0])
    _capi.check(_capi.lib.usgscsm_b200_image_to_ground_multi(
        handles, count, 2, line.ctypes.data, samp.ctypes.data, h1.ctypes.data, 0, 0.001,
        x.ctypes.data, y.ctypes.data, z.ctypes.data, None, None))
    ex, ey, ez, _, _ = gms[0].imageToGroundBatch(line[:2], samp[:2], 10.0)
    assert np.array_equal(x[:2], ex) and np.array_equal(z[:2], ez)
    assert _capi.lib.usgscsm_b200_ground_to_image_multi(handles, count, 0, None, None, None, 0.001,
                                                         None, None, None, None) == 0


def test_multi_rejects_bad_arguments():
    from usgscsm_b200 import _capi, synth
    gm = _models(synth.frame_state(), 1)[0]
    twice = (C.c_void_p * 2)(gm._h, gm._h)
    x = np.zeros(4)
    assert _capi.lib.usgscsm_b200_ground_to_image_multi(
        twice, 2, 4, x.ctypes.data, x.ctypes.data, x.ctypes.data, 0.001, x.ctypes.data, x.ctypes.data,
        None, None) == _capi.EINVAL
    assert b"twice" in _capi.lib.usgscsm_b200_last_error()
    assert _capi.lib.usgscsm_b200_ground_to_image_multi(
        None, 0, 4, x.ctypes.data, x.ctypes.data, x.ctypes.data, 0.001, x.ctypes.data, x.ctypes.data,
        None, None) == _capi.EINVAL


def test_concurrent_host_threads_on_different_models():
    """const calls are re-entrant in the reference; here several host threads drive different
    models (different kinds, same device) at once and each must get its own answers."""
    from usgscsm_b200 import synth
    states = [synth.linescan_state(n_lines=2048, n_samples=512, seed=2), synth.frame_state(),
              synth.pushframe_state(n_framelets=32, framelet_height=16, n_samples=512),
              synth.linescan_state(n_lines=1024, n_samples=256, seed=9)]
    gms = [_models(s, 1)[0] for s in states]
    rng = np.random.default_rng(4)
    work, expect = [], []
    for gm in gms:
        sz = gm.getImageSize()
        n = 600_000
        line, samp = rng.uniform(0, sz.line, n), rng.uniform(0, sz.samp, n)
        h = rng.uniform(-200, 200, n)
        work.append((line, samp, h))
        x, y, z, a, st = gm.imageToGroundBatch(line, samp, h)
        expect.append((x, y, z, st) + tuple(gm.groundToImageBatch(x, y, z)[:2]))
    results = [None] * len(gms)

    def run(k):
        out = []
        for _ in range(3):
            line, samp, h = work[k]
            x, y, z, a, st = gms[k].imageToGroundBatch(line, samp, h)
            l, s, _, _ = gms[k].groundToImageBatch(x, y, z)
            out.append((x, y, z, st, l, s))
        results[k] = out

    threads = [threading.Thread(target=run, args=(k,)) for k in range(len(gms))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for k in range(len(gms)):
        for got in results[k]:
            for u, v in zip(got, expect[k]):
                assert np.array_equal(u, v, equal_nan=True)


def test_concurrent_host_threads_on_the_same_model():
    """The reference's const per-point calls are re-entrant on ONE model object (SURVEY.md §5,
    include/usgscsm/UsgsAstroLsSensorModel.h:910-914).  Host-pointer calls on one handle share
    the handle's pipeline slots; they are serialised by a per-model mutex, so threads
    hammering the same handle with DIFFERENT inputs -- batches and single points -- must each
    get their own answers."""
    from usgscsm_b200 import synth
    gm = _models(synth.linescan_state(n_lines=4096, n_samples=1024, seed=5), 1)[0]
    rng = np.random.default_rng(9)
    nthreads = 6
    work, expect = [], []
    for k in range(nthreads):
        n = 5_000_000 if k == 0 else (1 if k == 1 else 300_000 + 1000 * k)   # several chunks / one point / one chunk
        line, samp = rng.uniform(0, 4096, n), rng.uniform(0, 1024, n)
        h = rng.uniform(-300, 300, n)
        x, y, z, a, st = gm.imageToGroundBatch(line, samp, h)
        work.append((line, samp, h))
        expect.append((x, y, z, st) + tuple(gm.groundToImageBatch(x, y, z)[:2]))
    results = [None] * nthreads

    def run(k):
        out = []
        for _ in range(4 if k else 2):
            line, samp, h = work[k]
            x, y, z, a, st = gm.imageToGroundBatch(line, samp, h)
            l, s, _, _ = gm.groundToImageBatch(x, y, z)
            out.append((x, y, z, st, l, s))
        results[k] = out

    threads = [threading.Thread(target=run, args=(k,)) for k in range(nthreads)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for k in range(nthreads):
        for got in results[k]:
            for u, v in zip(got, expect[k]):
                assert np.array_equal(u, v, equal_nan=True)


@pytest.mark.parametrize("kind", ["frame", "pushframe", "linescan", "sar"])
def test_many_images_per_call(kind):
    """usgscsm_b200_*_images_dev: several images of one kind in one call (kernels forked over
    internal streams) give exactly what one call per image gives."""
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import synth
    make = {"frame": lambda k: synth.frame_state(n_lines=300 + 20 * k, n_samples=256, seed=k),
            "pushframe": lambda k: synth.pushframe_state(n_framelets=16 + k, framelet_height=16, n_samples=256, seed=k),
            "linescan": lambda k: synth.linescan_state(n_lines=400 + 30 * k, n_samples=256, seed=40 + k),
            "sar": lambda k: synth.sar_state(n_lines=2000 + 100 * k, n_samples=300)}[kind]
    plugin = ub.UsgsAstroPlugin()
    models = [plugin.constructModelFromState(make(k)) for k in range(6)]
    dev = torch.device("cuda:0")
    sizes = [(int(m.getImageSize().line), int(m.getImageSize().samp)) for m in models]
    counts = [a * b for a, b in sizes]
    n = sum(counts)
    h = torch.linspace(-200.0, 300.0, n, dtype=torch.float64, device=dev)
    x, y, z, a, st = ub.imageToGroundGridImages(models, 0.5, 0.5, 1.0, 1.0, [s[0] for s in sizes],
                                                [s[1] for s in sizes], h, achieved=True)
    line, samp, a2, st2 = ub.groundToImageImages(models, counts, x, y, z, achieved=True)
    rows = torch.cat([(torch.arange(nl, device=dev, dtype=torch.float64) + 0.5).repeat_interleave(ns) for nl, ns in sizes])
    cols = torch.cat([(torch.arange(ns, device=dev, dtype=torch.float64) + 0.5).repeat(nl) for nl, ns in sizes])
    xe, ye, ze, ae, ste = ub.imageToGroundImages(models, counts, rows, cols, h, achieved=True)
    o = 0
    for m, (nl, ns), c in zip(models, sizes, counts):
        gx, gy, gz, ga, gst = m.imageToGroundGrid(0.5, 0.5, 1.0, 1.0, nl, ns, h[o:o + c], achieved=True)
        for u, v in ((x, gx), (y, gy), (z, gz), (a, ga), (xe, gx), (ze, gz)):
            assert torch.equal(u[o:o + c], v) or torch.allclose(u[o:o + c], v, rtol=0, atol=0, equal_nan=True)
        assert torch.equal(st[o:o + c], gst)
        gl, gs, gach, gst2 = m.groundToImageBatch(gx, gy, gz)
        assert torch.allclose(line[o:o + c], gl, rtol=0, atol=0, equal_nan=True)
        assert torch.allclose(samp[o:o + c], gs, rtol=0, atol=0, equal_nan=True)
        assert torch.equal(st2[o:o + c], gst2)
        o += c
    torch.cuda.synchronize()
