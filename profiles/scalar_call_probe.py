import sys, time
sys.path.insert(0, "/root/repo")
import numpy as np
import usgscsm_b200 as ub
from usgscsm_b200 import synth, _capi
gm = ub.UsgsAstroPlugin().constructModelFromState(synth.linescan_state())
x, y, z, _, _ = gm.imageToGroundBatch(np.array([1000.5]), np.array([2000.5]), np.array([0.0]))
import ctypes as C
lib = _capi.lib
l = np.zeros(1); s = np.zeros(1); st = np.zeros(1, dtype=np.uint32)
for n in (1, 16, 256):
    xs, ys, zs = np.repeat(x, n), np.repeat(y, n), np.repeat(z, n)
    l = np.zeros(n); s = np.zeros(n); st = np.zeros(n, dtype=np.uint32)
    for _ in range(50):
        lib.usgscsm_b200_ground_to_image(gm._h, n, xs.ctypes.data, ys.ctypes.data, zs.ctypes.data, 0.001, l.ctypes.data, s.ctypes.data, None, st.ctypes.data)
    t = time.perf_counter()
    R = 2000
    for _ in range(R):
        lib.usgscsm_b200_ground_to_image(gm._h, n, xs.ctypes.data, ys.ctypes.data, zs.ctypes.data, 0.001, l.ctypes.data, s.ctypes.data, None, st.ctypes.data)
    dt = (time.perf_counter() - t) / R
    print("groundToImage host arrays n=%d: %.1f us per call" % (n, dt * 1e6))
    t = time.perf_counter()
    for _ in range(R):
        lib.usgscsm_b200_image_to_ground(gm._h, n, l.ctypes.data, s.ctypes.data, zs.ctypes.data, 0, 0.001, xs.ctypes.data, ys.ctypes.data, zs.ctypes.data, None, st.ctypes.data)
    dt = (time.perf_counter() - t) / R
    print("imageToGround host arrays n=%d: %.1f us per call" % (n, dt * 1e6))
