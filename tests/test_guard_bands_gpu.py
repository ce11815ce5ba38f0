"""Output arrays with sentinel bands on either side: the vectorised kernels (two points per
128-bit access in the frame kernels, their scalar tails, the grid kernels' row / column
stepping) must write exactly the n points they are given -- odd and even n, arrays on and off
the 16-byte boundary."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GUARD = 64
SENTINEL = -12345.678


def _banded(n, off, dev, dtype):
    import torch
    fill = SENTINEL if dtype.is_floating_point else 0x5A5A5A5
    buf = torch.full((n + 2 * GUARD + 1,), fill, dtype=dtype, device=dev)
    return buf, buf[GUARD + off:GUARD + off + n]


def _intact(buf, n, off):
    import torch
    fill = SENTINEL if buf.dtype.is_floating_point else 0x5A5A5A5
    lo, hi = buf[:GUARD + off], buf[GUARD + off + n:]
    return bool((lo == fill).all()) and bool((hi == fill).all())


@pytest.mark.parametrize("kind", ["frame", "frame_transverse", "pushframe", "linescan", "sar"])
def test_outputs_stay_inside_their_arrays(kind):
    import torch
    import usgscsm_b200 as ub
    import ctypes as C
    from usgscsm_b200 import _capi, synth
    dev = torch.device("cuda:0")
    T = (0.0, 1.0, 0.0, 1e-6, 2e-6, -1e-6, 1e-8, 0.0, 2e-8, 0.0, 0.0, 0.0, 1.0, -1e-6, 1e-6, 2e-6, 0.0, 1e-8, 0.0, -2e-8)
    state = {"frame": lambda: synth.frame_state(n_lines=300, n_samples=258),
             "frame_transverse": lambda: synth.frame_state(n_lines=300, n_samples=258, distortion_type=synth.TRANSVERSE,
                                                           dist_coeffs=T),
             "pushframe": lambda: synth.pushframe_state(n_framelets=20, framelet_height=16, n_samples=258),
             "linescan": lambda: synth.linescan_state(n_lines=320, n_samples=258),
             "sar": lambda: synth.sar_state(n_lines=320, n_samples=258)}[kind]()
    gm = ub.UsgsAstroPlugin().constructModelFromState(state)
    sz = gm.getImageSize()
    for nl, ns in ((min(int(sz.line), 300), 258), (7, 257), (1, 1), (3, 2)):
        n = nl * ns
        for off in (0, 1):
            h = torch.linspace(-100.0, 100.0, n + 1, dtype=torch.float64, device=dev)[off:off + n]
            bufs = [_banded(n, off, dev, torch.float64) for _ in range(4)] + [_banded(n, off, dev, torch.int32)]
            out = tuple(v for _, v in bufs)
            gm.imageToGroundGrid(0.5, 0.5, 1.0, 1.0, nl, ns, h, achieved=True, out=out)
            torch.cuda.synchronize()
            assert all(_intact(b, n, off) for b, _ in bufs), (kind, "grid", nl, ns, off)
            assert bool((out[0] != SENTINEL).all()) and bool((out[4] != 0x5A5A5A5).all())
            x, y, z = out[0].clone(), out[1].clone(), out[2].clone()
            # explicit pixels and groundToImage, odd / even counts from both ends
            for m, o2 in ((n, off), (max(1, n - 1), 1 - off)):
                xs, ys, zs = x[:m], y[:m], z[:m]
                g = [_banded(m, o2, dev, torch.float64) for _ in range(3)] + [_banded(m, o2, dev, torch.int32)]
                gm.groundToImageBatch(xs, ys, zs, achieved=True, out=tuple(v for _, v in g))
                torch.cuda.synchronize()
                assert all(_intact(b, m, o2) for b, _ in g), (kind, "g2i", m, o2)
                line, samp = g[0][1], g[1][1]
                e = [_banded(m, o2, dev, torch.float64) for _ in range(4)] + [_banded(m, o2, dev, torch.int32)]
                _capi.check(_capi.lib.usgscsm_b200_image_to_ground_dev(
                    gm._h, m, line.data_ptr(), samp.data_ptr(), h[:m].data_ptr(), 1, 0.001, e[0][1].data_ptr(),
                    e[1][1].data_ptr(), e[2][1].data_ptr(), e[3][1].data_ptr(), e[4][1].data_ptr(),
                    C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
                torch.cuda.synchronize()
                assert all(_intact(b, m, o2) for b, _ in e), (kind, "i2g", m, o2)
                ok = torch.isfinite(e[0][1]) & torch.isfinite(xs)
                assert bool(ok.any())
                assert float((e[0][1][ok] - xs[ok]).abs().max()) < 1.0
