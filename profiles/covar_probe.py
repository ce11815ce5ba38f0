"""groundToImage with covariance (22 searches per point) on 8 Mi points of the C2 model: wall time
per call; under `ncu --metrics gpu__time_duration.sum` the launch list shows where it goes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import usgscsm_b200 as ub
from usgscsm_b200 import _capi, synth

dev = torch.device("cuda:0")
gm = ub.UsgsAstroPlugin().constructModelFromState(synth.linescan_state())
n = 1 << 23
nl, ns = 1664, 5042
h = torch.zeros(nl * ns, dtype=torch.float64, device=dev)
x, y, z, _, _ = gm.imageToGroundGrid(20000.5, 10.5, 1.0, 1.0, nl, ns, h, achieved=False)
x, y, z = x[:n].contiguous(), y[:n].contiguous(), z[:n].contiguous()
gcov = torch.eye(3, dtype=torch.float64, device=dev).reshape(1, 9).repeat(n, 1).contiguous()
icov = torch.empty(n * 4, dtype=torch.float64, device=dev)
o = [torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3)]
st = torch.empty(n, dtype=torch.int32, device=dev)
p = lambda t: t.data_ptr()
stream = torch.cuda.current_stream(dev).cuda_stream
npar = _capi.lib.usgscsm_b200_model_num_parameters(gm._h)
allp = torch.empty(n * npar * 2, dtype=torch.float64, device=dev)


def covar():
    _capi.check(_capi.lib.usgscsm_b200_ground_to_image_covar_dev(
        gm._h, n, p(x), p(y), p(z), p(gcov), 0.001, p(o[0]), p(o[1]), p(icov), p(o[2]), p(st), stream))


def partials():
    _capi.check(_capi.lib.usgscsm_b200_all_sensor_partials_dev(
        gm._h, npar, None, n, p(x), p(y), p(z), None, None, 0.001, p(allp), p(o[0]), p(st), stream))


for name, fn in (("covar", covar), ("all sensor partials", partials)):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3):
        fn()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 3
    print("%s: %.2f ms for %d points = %.0f Mpt/s" % (name, ms, n, n / ms / 1e3))
