"""Why a C5 pass (many band launches back to back, one model each) takes longer than the sum of
its kernels: per-launch CUDA events + the span of the whole pass, on one GPU."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import usgscsm_b200 as ub
from usgscsm_b200 import _capi, synth

dev = torch.device("cuda:0")
images, lines, ns = int(sys.argv[1]) if len(sys.argv) > 1 else 24, 24684, 5064
nb = lines * ns
plugin = ub.UsgsAstroPlugin()
models = [plugin.constructModelFromState(synth.linescan_state(n_lines=49368, n_samples=ns, seed=20261105 + i, polar=True))
          for i in range(images)]
h = torch.zeros(nb, dtype=torch.float64, device=dev)
gx, gy, gz = (torch.empty(images * nb, dtype=torch.float64, device=dev) for _ in range(3))
ol, os_ = (torch.empty(4 * nb, dtype=torch.float64, device=dev) for _ in range(2))
ost = torch.empty(4 * nb, dtype=torch.int32, device=dev)
stg = torch.empty(nb, dtype=torch.int32, device=dev)
stream = torch.cuda.current_stream(dev).cuda_stream
p = lambda t: t.data_ptr()
lib = _capi.lib
for k, m in enumerate(models):
    _capi.check(lib.usgscsm_b200_image_to_ground_grid_dev(m._h, 0.5, 0.5, 1.0, 1.0, lines, ns, p(h), 1, 0.001,
                                                          p(gx) + 8 * k * nb, p(gy) + 8 * k * nb, p(gz) + 8 * k * nb,
                                                          None, p(stg), stream))
torch.cuda.synchronize()


def launch(k, model_index):
    r = k % 4
    _capi.check(lib.usgscsm_b200_ground_to_image_dev(models[model_index]._h, nb, p(gx) + 8 * k * nb, p(gy) + 8 * k * nb,
                                                     p(gz) + 8 * k * nb, 0.001, p(ol) + 8 * r * nb, p(os_) + 8 * r * nb,
                                                     None, p(ost) + 4 * r * nb, stream))


for label, pick in (("one model each", lambda k: k), ("same model, different inputs", lambda k: 0)):
    for rep in range(2):
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(images)]
        t0 = time.perf_counter()
        for k in range(images):
            evs[k][0].record()
            launch(k, pick(k))
            evs[k][1].record()
        host = time.perf_counter() - t0
        torch.cuda.synchronize()
        per = [a.elapsed_time(b) for a, b in evs]
        span = evs[0][0].elapsed_time(evs[-1][1])
        gaps = [evs[k][1].elapsed_time(evs[k + 1][0]) for k in range(images - 1)]
        print("%s, rep %d: host enqueue %.1f ms, span %.1f ms, sum of kernels %.1f ms, per-launch min/med/max %.2f/%.2f/%.2f, "
              "gaps med/max %.3f/%.3f ms" % (label, rep, host * 1e3, span, sum(per), min(per), sorted(per)[len(per) // 2],
                                             max(per), sorted(gaps)[len(gaps) // 2], max(gaps)), flush=True)
        print("   first 8 launches:", " ".join("%.2f" % v for v in per[:8]), flush=True)
        if label.startswith("one") and rep == 0:
            for k in range(images):
                launch(k, k)
                r = k % 4
                st = ost[r * nb:(r + 1) * nb]
                ev = ((st >> _capi.ST_EVALS_SHIFT) & 0xff)
                d = models[k].lsDerived()
                print("   image %2d: %.2f ms, evaluations mean %.3f max %d, flags %d, halfSwath %.1f gsd %.4f seed used %d"
                      % (k, per[k], float(ev.double().mean()), int(ev.max()), int(((st & _capi.ST_REFERENCE_MASK) != 0).sum()),
                         d[2], d[0], int(d[4])), flush=True)
