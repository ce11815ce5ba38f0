"""Cost of cudaHostRegister / cudaHostUnregister on a pageable numpy array against the
transfer-rate gain (decides whether the host-pointer entry points should pin the caller's pages
for the duration of a call).  Run on the GPU box."""
import time

import numpy as np
import torch

rt = torch.cuda.cudart()
torch.cuda.init()
dev = torch.empty(1 << 28, dtype=torch.float64, device="cuda")   # 2 GiB
for mb in (64, 512, 2048):
    a = np.random.default_rng(0).random(mb * 1024 * 1024 // 8)
    t = time.perf_counter()
    dev[:a.size].copy_(torch.from_numpy(a))
    torch.cuda.synchronize()
    t_page = time.perf_counter() - t
    t = time.perf_counter()
    rc = rt.cudaHostRegister(a.ctypes.data, a.nbytes, 0)
    t_reg = time.perf_counter() - t
    t = time.perf_counter()
    dev[:a.size].copy_(torch.from_numpy(a), non_blocking=True)
    torch.cuda.synchronize()
    t_pin = time.perf_counter() - t
    t = time.perf_counter()
    rt.cudaHostUnregister(a.ctypes.data)
    t_unreg = time.perf_counter() - t
    print("%5d MB: pageable copy %.1f GB/s | register %.1f ms (%.1f GB/s) rc=%s, pinned copy %.1f GB/s, unregister %.1f ms"
          % (mb, a.nbytes / t_page / 1e9, t_reg * 1e3, a.nbytes / t_reg / 1e9, rc, a.nbytes / t_pin / 1e9, t_unreg * 1e3))
