// work_queue.cuh -- dynamic distribution of the points of one launch over the resident warps.
//
// The kernels run persistent grids (blocks = SMs x resident blocks) so that table staging is
// paid once per block.  With a static grid-stride split the warps the schedulers favour finish
// their share early and the SM drains at reduced occupancy (ncu, line-scan groundToImage:
// 24.9 of 32 resident warps active on average); per-point iteration counts that differ
// (secant / Brent searches) make it worse.  Here each warp takes `grab` consecutive points
// at a time from a counter in global memory; a warp still touches 32 consecutive doubles per
// array per pass (two full 128-byte lines).
#ifndef USGSCSM_B200_WORK_QUEUE_CUH_
#define USGSCSM_B200_WORK_QUEUE_CUH_

#include <cuda_runtime.h>

namespace ub200 {

constexpr int kDefaultGrab = 512;  // points per grab (a multiple of 32)

// body(i) is called once for every i in [0, n); all 32 lanes of a warp must enter together.
template <class F>
__device__ __forceinline__ void forEachPointDynamic(long n, unsigned long long *__restrict__ next,
                                                    int grab, F &&body) {
  const unsigned lane = threadIdx.x & 31u;
  for (;;) {
    __syncwarp();
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(next, (unsigned long long)grab);
    base = __shfl_sync(0xffffffffu, base, 0);
    if ((long)base >= n) break;
#pragma unroll 1
    for (int k = 0; k < grab; k += 32) {
      const long i = (long)base + k + lane;
      if (i >= n) break;
      body(i);
    }
  }
}

// Host side: the counter of one launch.  Stream-ordered scratch, so launches in flight on
// different streams never share a counter; zeroed on the launch stream before the kernel.
struct WorkCounter {
  unsigned long long *ptr = nullptr;
  cudaStream_t stream = nullptr;
  // (n, the launch's point count, is unused: a counter-free static split for small launches cost
  // the search kernel registers -- 0.8 % of its throughput -- for 5 us of a scalar call)
  cudaError_t acquire(cudaStream_t s, long /*n*/ = -1) {
    stream = s;
    cudaError_t e = cudaMallocAsync(&ptr, sizeof(*ptr), s);
    if (e != cudaSuccess) return e;
    return cudaMemsetAsync(ptr, 0, sizeof(*ptr), s);
  }
  // call after the kernel launch; returns the launch error if there was one
  cudaError_t release() {
    cudaError_t e = cudaGetLastError();
    cudaError_t e2 = ptr ? cudaFreeAsync(ptr, stream) : cudaSuccess;
    ptr = nullptr;
    return e != cudaSuccess ? e : e2;
  }
};

int grabSize();  // USGSCSM_B200_GRAB or kDefaultGrab, rounded up to a multiple of 32 (capi.cu)
// Small launches: one warp's worth per grab, so that the few points spread over the block's warps
// instead of queueing behind one of them (256 points: 65 -> 35 us per groundToImage call).
inline int grabFor(long n) { return n <= 8192 ? 32 : grabSize(); }

}  // namespace ub200
#endif
