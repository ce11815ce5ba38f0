"""Throughput on the reference's own HiRISE state (tests/golden/ref_data/gxp_model_file.sup: 40000 x
20000 pixels, 40 003 quaternions = a 1.28 MB attitude table that does not fit shared memory):
dense-grid imageToGround and groundToImage of a band of 8000 lines."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import usgscsm_b200 as ub

dev = torch.device("cuda:0")
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "ref_data",
                    "gxp_model_file.sup")
gm = ub.UsgsAstroPlugin().constructModelFromState(open(path, "rb").read().decode("latin-1"))
sz = gm.getImageSize()
print("image", int(sz.line), "x", int(sz.samp), gm.getModelName())
nl, ns = 8000, int(sz.samp)
n = nl * ns
h = torch.zeros(n, dtype=torch.float64, device=dev)


def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


out = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3)) + (None, torch.empty(n, dtype=torch.int32, device=dev))
ms = timed(lambda: gm.imageToGroundGrid(16000.5, 0.5, 1.0, 1.0, nl, ns, h, out=out))
print("imageToGround grid: %.2f ms, %.1f Gpt/s" % (ms, n / ms / 1e6))
x, y, z = out[0], out[1], out[2]
o2 = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2)) + (None, torch.empty(n, dtype=torch.int32, device=dev))
ms = timed(lambda: gm.groundToImageBatch(x, y, z, achieved=False, out=o2))
ev = float(((o2[3] >> 16) & 0xff).double().mean()) if hasattr(ub, "_capi") else 0.0
print("groundToImage: %.2f ms, %.2f Gpt/s" % (ms, n / ms / 1e6))
rows = torch.arange(nl, device=dev, dtype=torch.float64) + 16000.5
print("round trip max |dline| %.2e" % float((o2[0].view(nl, ns) - rows[:, None]).abs().max()))
