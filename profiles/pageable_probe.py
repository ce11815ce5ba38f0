"""End-to-end rate of the host-pointer entry points with PAGEABLE numpy arrays, with and without
the one-thread-per-slot pipeline (USGSCSM_B200_PIPELINE_THREADS), against pinned arrays."""
import os
import subprocess
import sys
import time

import numpy as np

if len(sys.argv) > 1:
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import synth
    gm = ub.UsgsAstroPlugin().constructModelFromState(synth.linescan_state())
    nl, ns = 8192, 5064
    line = np.repeat(np.arange(nl) + 0.5, ns)
    samp = np.tile(np.arange(ns) + 0.5, nl)
    h = synth.synthetic_dem(line, samp)
    n = line.size
    x, y, z, _, _ = gm.imageToGroundBatch(line, samp, h, 1e-3)
    from usgscsm_b200 import _capi
    ol, os_ = np.zeros(n), np.zeros(n)
    st = np.zeros(n, dtype=np.uint32)

    def call():
        _capi.check(_capi.lib.usgscsm_b200_ground_to_image(
            gm._h, n, x.ctypes.data, y.ctypes.data, z.ctypes.data, 1e-3, ol.ctypes.data,
            os_.ctypes.data, None, st.ctypes.data))
    call()
    t = time.perf_counter()
    for _ in range(3):
        call()
    dt = (time.perf_counter() - t) / 3
    print("pageable numpy arrays, USGSCSM_B200_PIPELINE_THREADS=%s: groundToImage %d points, %.1f ms, %.0f Mpoints/s"
          % (os.environ.get("USGSCSM_B200_PIPELINE_THREADS", "1"), n, dt * 1e3, n / dt / 1e6), flush=True)
    print("round trip max |dline| %.3g px" % float(np.abs(ol - line).max()))
else:
    for v, slots, chunk in (("1", "3", str(1 << 22)), ("0", "3", str(1 << 22)), ("1", "6", str(1 << 20)),
                            ("1", "8", str(1 << 21))):
        env = dict(os.environ, USGSCSM_B200_PIPELINE_THREADS=v, USGSCSM_B200_PIPELINE_SLOTS=slots,
                   USGSCSM_B200_PIPELINE_CHUNK=chunk)
        print("slots %s, chunk %s points:" % (slots, chunk), flush=True)
        subprocess.run([sys.executable, __file__, "child"], env=env, check=True)
