"""16 images of the C3 size (4096 x 4096) per call: one call per image on one stream, the
many-images entry points (kernels forked over side streams), and ONE image 16 times as tall as the
upper bound a single launch over all states could reach.  Frame and push-frame, both directions."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import usgscsm_b200 as ub
from usgscsm_b200 import synth

dev = torch.device("cuda:0")
plugin = ub.UsgsAstroPlugin()
K = 16


def timed(fn, reps=5):
    fn(); fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return sorted(ms)[len(ms) // 2]


for name, make, tall in (
        ("frame radial", lambda k: synth.frame_state(n_lines=4096, n_samples=4096, seed=5 + k),
         lambda: synth.frame_state(n_lines=4096 * K, n_samples=4096)),
        ("push-frame radial", lambda k: synth.pushframe_state(n_framelets=256, framelet_height=16, n_samples=4096, seed=3 + k),
         lambda: synth.pushframe_state(n_framelets=256 * K, framelet_height=16, n_samples=4096))):
    models = [plugin.constructModelFromState(make(k)) for k in range(K)]
    nl, ns = int(models[0].getImageSize().line), int(models[0].getImageSize().samp)
    n1 = nl * ns
    n = n1 * K
    h = torch.zeros(n, dtype=torch.float64, device=dev)
    o1 = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3)) + (None, torch.empty(n, dtype=torch.int32, device=dev))
    o2 = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2)) + (None, torch.empty(n, dtype=torch.int32, device=dev))

    def views(o, k):
        return tuple(None if t is None else t[k * n1:(k + 1) * n1] for t in o)

    def i2g_each():
        for k, m in enumerate(models):
            m.imageToGroundGrid(0.5, 0.5, 1.0, 1.0, nl, ns, h[k * n1:(k + 1) * n1], out=views(o1, k))

    def i2g_images():
        ub.imageToGroundGridImages(models, 0.5, 0.5, 1.0, 1.0, [nl] * K, [ns] * K, h, out=o1)

    def g2i_each():
        for k, m in enumerate(models):
            x, y, z, _, _ = views(o1, k)
            m.groundToImageBatch(x, y, z, achieved=False, out=views(o2, k))

    def g2i_images():
        ub.groundToImageImages(models, [n1] * K, o1[0], o1[1], o1[2], out=o2)

    i2g_images()
    res = {"i2g one call per image": timed(i2g_each), "i2g images call": timed(i2g_images),
           "g2i one call per image": timed(g2i_each), "g2i images call": timed(g2i_images)}
    big = plugin.constructModelFromState(tall())
    bl, bs = int(big.getImageSize().line), int(big.getImageSize().samp)
    nb = bl * bs
    hb = torch.zeros(nb, dtype=torch.float64, device=dev)
    b1 = tuple(torch.empty(nb, dtype=torch.float64, device=dev) for _ in range(3)) + (None, torch.empty(nb, dtype=torch.int32, device=dev))
    b2 = tuple(torch.empty(nb, dtype=torch.float64, device=dev) for _ in range(2)) + (None, torch.empty(nb, dtype=torch.int32, device=dev))
    res["i2g ONE tall image (%d pts)" % nb] = timed(lambda: big.imageToGroundGrid(0.5, 0.5, 1.0, 1.0, bl, bs, hb, out=b1)) * n / nb
    res["g2i ONE tall image"] = timed(lambda: big.groundToImageBatch(b1[0], b1[1], b1[2], achieved=False, out=b2)) * n / nb
    print(name, "(%d x %d x %d = %.0f M points)" % (K, nl, ns, n / 1e6))
    for k, ms in res.items():
        bytes_pt = 36 if k.startswith("i2g") else 44
        print("   %-34s %7.3f ms  %7.1f Gpt/s  %5.2f TB/s" % (k, ms, n / ms / 1e6, bytes_pt * n / ms / 1e9))
    del models, big, h, hb, o1, o2, b1, b2
    torch.cuda.empty_cache()
