"""usgscsm_b200_image_to_ground_grid_static_dev: the dense-grid imageToGround of the line-scan
model with its row / column tables in a file-scope device array (no cudaMalloc of any kind),
against usgscsm_b200_image_to_ground_grid_dev through the C-ABI.  The two run the same kernels, so
the comparison is bit for bit (the bar of tests/test_ls_parity_gpu.py's grid / explicit test).
Two calls in a row on different grids: the array keeps the first call's tables, the second must
not see them.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _grid_call(fn, gm, line0, samp0, dline, dsamp, nl, ns, h, prec, stream=None):
    import torch
    from usgscsm_b200 import _capi
    n = nl * ns
    dev = h.device
    x, y, z, a = (torch.full((n,), -7.0, dtype=torch.float64, device=dev) for _ in range(4))
    st = torch.full((n,), 0x7fffffff, dtype=torch.int32, device=dev)
    s = torch.cuda.current_stream(dev).cuda_stream if stream is None else stream
    rc = fn(gm._h, line0, samp0, dline, dsamp, nl, ns, h.data_ptr(), 1, prec, x.data_ptr(), y.data_ptr(),
            z.data_ptr(), a.data_ptr(), st.data_ptr(), C.c_void_p(s))
    _capi.check(rc)
    return x, y, z, a, st


def _same(u, v):
    import torch
    torch.cuda.synchronize()
    for p, q in zip(u[:4], v[:4]):
        assert np.array_equal(p.cpu().numpy().view(np.uint64), q.cpu().numpy().view(np.uint64))
    assert np.array_equal(u[4].cpu().numpy(), v[4].cpu().numpy())


@pytest.mark.parametrize("adjusted", [False, True])
def test_static_scratch_grid_matches_the_allocating_entry_point(adjusted):
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, synth
    dev = torch.device("cuda:0")
    gm = ub.UsgsAstroPlugin().constructModelFromState(synth.linescan_state(n_lines=4096, n_samples=1024, seed=21))
    if adjusted:
        for i, v in {0: 2.0, 4: -0.01, 7: 1e-4, 11: 2e-6, 15: 0.02}.items():
            gm.setParameterValue(i, v)
    lib = _capi.lib
    # first call: the larger grid; second: fewer, other rows and columns (stale tables would show)
    grids = [(0.5, 0.5, 1.0, 1.0, 4096, 1024), (100.3, 7.7, 0.1, 1.3, 777, 311)]
    hs = [torch.linspace(-400.0, 600.0, nl * ns, dtype=torch.float64, device=dev) for *_, nl, ns in grids]
    want = [_grid_call(lib.usgscsm_b200_image_to_ground_grid_dev, gm, *g, h, 0.001) for g, h in zip(grids, hs)]
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]
    outs = [tuple(torch.full((g[4] * g[5],), -7.0, dtype=torch.float64, device=dev) for _ in range(4)) +
            (torch.full((g[4] * g[5],), 0x7fffffff, dtype=torch.int32, device=dev),) for g in grids]
    torch.cuda.synchronize()
    free_ready = torch.cuda.mem_get_info(dev)[0]
    s = torch.cuda.current_stream(dev).cuda_stream
    for g, h, o in zip(grids, hs, outs):        # two calls in a row, nothing between them
        _capi.check(lib.usgscsm_b200_image_to_ground_grid_static_dev(
            gm._h, g[0], g[1], g[2], g[3], g[4], g[5], h.data_ptr(), 1, 0.001, o[0].data_ptr(), o[1].data_ptr(),
            o[2].data_ptr(), o[3].data_ptr(), o[4].data_ptr(), C.c_void_p(s)))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] >= free_ready, (free_before, free_ready)   # nothing was allocated
    for w, o in zip(want, outs):
        _same(w, o)
    # and once more in the other order: the large grid after the small one
    again = _grid_call(lib.usgscsm_b200_image_to_ground_grid_static_dev, gm, *grids[0], hs[0], 0.001)
    _same(want[0], again)
    # sanity: real coordinates, not the fill value
    assert float(want[0][0].abs().min()) > 1.0 or float(want[0][1].abs().min()) > 1.0


def test_static_scratch_limits_are_errors():
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, synth
    dev = torch.device("cuda:0")
    lib = _capi.lib
    gm = ub.UsgsAstroPlugin().constructModelFromState(synth.linescan_state(n_lines=4096, n_samples=1024, seed=21))
    h = torch.zeros(16, dtype=torch.float64, device=dev)
    o = torch.zeros(16, dtype=torch.float64, device=dev)
    null = C.c_void_p(0)
    # 21 doubles per row: 99865 rows need one double more than the 2 Mi the array holds
    rc = lib.usgscsm_b200_image_to_ground_grid_static_dev(gm._h, 0.5, 0.5, 1.0, 1.0, 99865, 1, h.data_ptr(), 0, 0.001,
                                                          o.data_ptr(), o.data_ptr(), o.data_ptr(), None, None, null)
    assert rc == _capi.EINVAL and b"too large" in lib.usgscsm_b200_last_error()
    # the largest grid of one column that fits is accepted (height broadcast, outputs of that size)
    n = 99864
    big = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3))
    rc = lib.usgscsm_b200_image_to_ground_grid_static_dev(gm._h, 0.5, 0.5, 0.01, 1.0, n, 1, h.data_ptr(), 0, 0.001,
                                                          big[0].data_ptr(), big[1].data_ptr(), big[2].data_ptr(),
                                                          None, None, null)
    assert rc == _capi.OK
    torch.cuda.synchronize()
    assert bool(torch.isfinite(big[0]).all())
    # empty grids are fine, other models are refused
    assert lib.usgscsm_b200_image_to_ground_grid_static_dev(gm._h, 0.5, 0.5, 1.0, 1.0, 0, 5, None, 0, 0.001, None, None,
                                                            None, None, None, null) == _capi.OK
    fr = ub.UsgsAstroPlugin().constructModelFromState(synth.frame_state(n_lines=64, n_samples=64))
    rc = lib.usgscsm_b200_image_to_ground_grid_static_dev(fr._h, 0.5, 0.5, 1.0, 1.0, 4, 4, h.data_ptr(), 1, 0.001,
                                                          o.data_ptr(), o.data_ptr(), o.data_ptr(), None, None, null)
    assert rc == _capi.EUNSUPPORTED
