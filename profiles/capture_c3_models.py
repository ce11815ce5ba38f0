"""Launch the push-frame and frame kernels of config C3 (4096 x 4096, DEM heights) twice each so
that one `ncu --set full` run captures them:
    python profiles/capture_c3_models.py                      # must exit 0 first
    ncu --set full --clock-control none --import-source on -k regex:'^(pf|fr)(Ground|Image)' \
        -o gpurun_out/r2_c3_full python profiles/capture_c3_models.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import usgscsm_b200 as ub  # noqa: E402
from usgscsm_b200 import synth  # noqa: E402

dev = torch.device("cuda", 0)
cases = [
    ("pushframe", synth.pushframe_state(n_framelets=256, framelet_height=16, n_samples=4096)),
    ("frame_radial", synth.frame_state(n_lines=4096, n_samples=4096)),
]
for name, state in cases:
    gm = ub.UsgsAstroPlugin(device=0).constructModelFromState(state)
    sz = gm.getImageSize()
    nl, ns = int(sz.line), int(sz.samp)
    rows = torch.arange(nl, device=dev, dtype=torch.float64) + 0.5
    cols = torch.arange(ns, device=dev, dtype=torch.float64) + 0.5
    h = (500.0 * torch.sin(2 * np.pi * rows / 4096.0)[:, None] *
         torch.cos(2 * np.pi * cols / 1024.0)[None, :]).reshape(-1).contiguous()
    for _ in range(2):
        x, y, z, _, st = gm.imageToGroundGrid(0.5, 0.5, 1.0, 1.0, nl, ns, h, 0.001)
        line, samp, _, st2 = gm.groundToImageBatch(x, y, z, 0.001)
    torch.cuda.synchronize()
    print(name, nl * ns, "points, round trip",
          float((line.view(nl, ns) - rows[:, None]).abs().max()),
          float((samp.view(nl, ns) - cols[None, :]).abs().max()))
