"""End-to-end rate of the host-pointer groundToImage with PAGEABLE numpy arrays (what a numpy
caller passes), round 2: the engine stages them through its own page-locked ring, one copier
thread per slot (csrc/pipeline.cuh).  Sweeps USGSCSM_B200_STAGE_THREADS x USGSCSM_B200_STAGE_CHUNK
and compares with the CUDA runtime's own staging (USGSCSM_B200_PIPELINE_THREADS=0) and with plain
host memcpy bandwidth of the box.  Outputs preallocated and touched (no page faults in the timed
region); a second line times the call with fresh numpy.empty outputs."""
import os
import subprocess
import sys
import time

import numpy as np

if len(sys.argv) > 1:
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import usgscsm_b200 as ub
    from usgscsm_b200 import synth, _capi
    gm = ub.UsgsAstroPlugin().constructModelFromState(synth.linescan_state())
    nl, ns = 12288, 5064
    line = np.repeat(np.arange(nl) + 0.5, ns)
    samp = np.tile(np.arange(ns) + 0.5, nl)
    h = synth.synthetic_dem(line, samp)
    n = line.size
    x, y, z, _, _ = gm.imageToGroundBatch(line, samp, h, 1e-3)
    ol, os_ = np.zeros(n), np.zeros(n)
    st = np.zeros(n, dtype=np.uint32)

    def call(ol, os_, st):
        _capi.check(_capi.lib.usgscsm_b200_ground_to_image(
            gm._h, n, x.ctypes.data, y.ctypes.data, z.ctypes.data, 1e-3, ol.ctypes.data,
            os_.ctypes.data, None, st.ctypes.data))
    call(ol, os_, st)
    t = time.perf_counter()
    for _ in range(3):
        call(ol, os_, st)
    dt = (time.perf_counter() - t) / 3
    t = time.perf_counter()
    call(np.empty(n), np.empty(n), np.empty(n, dtype=np.uint32))
    dt2 = time.perf_counter() - t
    print("  threads %s chunk %s runtime-staging-off %s: %d points, %.1f ms, %.0f Mpoints/s (%.1f GB/s of "
          "caller arrays); fresh output arrays: %.0f Mpoints/s; round trip %.3g px"
          % (os.environ.get("USGSCSM_B200_STAGE_THREADS", "8"), os.environ.get("USGSCSM_B200_STAGE_CHUNK", str(1 << 20)),
             os.environ.get("USGSCSM_B200_PIPELINE_THREADS", "1"), n, dt * 1e3, n / dt / 1e6, 44 * n / dt / 1e9,
             n / dt2 / 1e6, float(np.abs(ol - line).max())), flush=True)
else:
    a = np.ones(1 << 27)
    b = np.empty_like(a)
    b[:] = a
    t = time.perf_counter()
    b[:] = a
    print("host memcpy, one thread: %.1f GB/s (read + write counted once)" % (a.nbytes / (time.perf_counter() - t) / 1e9))
    del a, b
    print("CUDA runtime staging (round 1 path):", flush=True)
    subprocess.run([sys.executable, __file__, "child"], env=dict(os.environ, USGSCSM_B200_PIPELINE_THREADS="0"), check=True)
    print("own pinned ring:", flush=True)
    for threads in ("4", "8", "12", "16"):
        for chunk in (str(1 << 18), str(1 << 20)):
            env = dict(os.environ, USGSCSM_B200_STAGE_THREADS=threads, USGSCSM_B200_STAGE_CHUNK=chunk)
            subprocess.run([sys.executable, __file__, "child"], env=env, check=True)
