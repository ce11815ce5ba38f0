"""A/B of the SAR groundToImage kernel variants (USGSCSM_B200_SAR_REFILL / _PAD / _MINB / _INTERLEAVE) on the
C4 workload (unordered random ground points, 0.1 % without a Doppler bracket) and on an ordered
dense grid.  Each setting in its own process."""
import itertools
import os
import subprocess
import sys

if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import numpy as np
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import synth, _capi
    dev = torch.device("cuda:0")
    gm = ub.UsgsAstroPlugin().constructModelFromState(synth.sar_state())
    nl, ns = int(gm.getImageSize().line), int(gm.getImageSize().samp)
    n = 1 << 25
    gen = np.random.Generator(np.random.MT19937(4))
    line = torch.as_tensor(gen.uniform(0.0, nl, n), device=dev)
    samp = torch.as_tensor(gen.uniform(0.0, ns, n), device=dev)
    hh = torch.as_tensor(gen.uniform(-1000.0, 1000.0, n), device=dev)
    res = []
    for name, (l, s) in (("random", (line, samp)),
                         ("ordered", (torch.sort(line)[0], samp))):
        x, y, z, _, _ = gm.imageToGroundBatch(l, s, hh, 0.001, achieved=False)
        out = tuple(torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2)) + (
            None, torch.empty(n, dtype=torch.int32, device=dev))
        for _ in range(2):
            gm.groundToImageBatch(x, y, z, 0.001, achieved=False, out=out)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(3):
            gm.groundToImageBatch(x, y, z, 0.001, achieved=False, out=out)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 3
        err = float((out[0] - l).abs().max())
        res.append("%s %.0f Mpt/s (%.1f ms, round trip %.2g px)" % (name, n / ms / 1e3, ms, err))
    print("  ".join(res), flush=True)
else:
    # "quick": only the table layouts of the lockstep kernel
    quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
    for refill, pad, minb, inter in itertools.product("0" if quick else "01", "01", "32", "10"):
        if pad == "1" and inter == "1":
            continue
        env = dict(os.environ, USGSCSM_B200_SAR_REFILL=refill, USGSCSM_B200_SAR_PAD=pad, USGSCSM_B200_SAR_MINB=minb,
                   USGSCSM_B200_SAR_INTERLEAVE=inter)
        print("refill %s pad %s minb %s interleave %s:" % (refill, pad, minb, inter), end=" ", flush=True)
        subprocess.run([sys.executable, __file__, "child"], env=env)
