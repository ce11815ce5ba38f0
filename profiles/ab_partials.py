"""A/B of the fused line-scan computeAllSensorPartials kernel (register budgets, fused vs the
per-parameter launches) on the C2 model, 8 Mi points, unadjusted and adjusted model."""
import os
import subprocess
import sys

if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import synth
    dev = torch.device("cuda:0")
    gm = ub.UsgsAstroPlugin().constructModelFromState(synth.linescan_state(polar=True))
    nl, ns = 1656, 5064
    n = nl * ns
    h = torch.zeros(1, dtype=torch.float64, device=dev)
    x, y, z, _, _ = gm.imageToGroundGrid(20000.5, 0.5, 1.0, 1.0, nl, ns, h)
    out = []
    for label, params in (("unadjusted", {}), ("adjusted", {0: 2.0, 4: -0.01, 7: 2e-5, 15: 0.02})):
        for i, v in params.items():
            gm.setParameterValue(i, v)
        for _ in range(2):
            gm.computeAllSensorPartialsBatch(x, y, z)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(3):
            r = gm.computeAllSensorPartialsBatch(x, y, z)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 3
        out.append("%s %.0f Mpt/s (%.1f ms; %.2f G searches/s)" % (label, n / ms / 1e3, ms, 17 * n / ms / 1e6))
    print("  ".join(out), flush=True)
else:
    for fused, minb in (("0", "3"), ("1", "2"), ("1", "3"), ("1", "4")):
        env = dict(os.environ, USGSCSM_B200_LS_FUSED_PARTIALS=fused, USGSCSM_B200_LS_PARTIALS_MINB=minb)
        print("fused %s minb %s:" % (fused, minb), end=" ", flush=True)
        subprocess.run([sys.executable, __file__, "child"], env=env)
