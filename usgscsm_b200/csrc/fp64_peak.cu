// fp64_peak.cu -- measures the FP64 pipe this engine's kernels are bound by (the roofline
// denominator of bench.py; BASELINE.md section 3 / VERDICT r1 "a measured FP64 peak").
//
// One kernel per instruction kind, dependent-free: every thread carries kChains independent
// accumulators, so that each warp always has an instruction ready; all SMs, 4 blocks x 256
// threads per SM (the shape of the search kernels), timed with CUDA events over repeated
// launches until at least `min_ms` of kernel time have accumulated.
//   op 0: DFMA  (2 flops / instruction) -- the datasheet "FP64 TFLOP/s" figure
//   op 1: DADD  (1 flop)
//   op 2: DMUL  (1 flop)
//   op 3: DADD + DMUL alternating -- the mix bit-exact parity allows (-fmad=false: the
//         reference's x86-64 build has no FMA contraction, so neither have the kernels)
// Result: FP64 thread-instructions per second.
#include <cuda_runtime.h>

#include "../../include/usgscsm_b200.h"
#include "capi_common.h"

namespace ub200 {
namespace {

constexpr int kChains = 8;
constexpr int kInner = 512;

template <int OP>
__global__ void __launch_bounds__(256, 4) fp64PeakKernel(double seed, int outer, double *sink) {
  double a[kChains];
#pragma unroll
  for (int k = 0; k < kChains; k++) a[k] = seed + (double)(threadIdx.x + k);
  const double b = 1.0000000001, c = 1e-9, bi = 0.9999999999;
  for (int o = 0; o < outer; o++) {
#pragma unroll 4
    for (int i = 0; i < kInner; i += 2) {
#pragma unroll
      for (int k = 0; k < kChains; k++) {
        if (OP == 0) {
          a[k] = fma(a[k], b, c);
          a[k] = fma(a[k], bi, c);
        } else if (OP == 1) {
          a[k] = __dadd_rn(a[k], c);
          a[k] = __dadd_rn(a[k], b);
        } else if (OP == 2) {
          a[k] = __dmul_rn(a[k], b);
          a[k] = __dmul_rn(a[k], bi);
        } else {
          a[k] = __dadd_rn(a[k], c);
          a[k] = __dmul_rn(a[k], bi);
        }
      }
    }
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < kChains; k++) s += a[k];
  if (s == 12345.6789) sink[0] = s;  // never true: keeps the chains alive
}

template <int OP>
cudaError_t runPeak(int blocks, int outer, double *sink, cudaStream_t stream) {
  fp64PeakKernel<OP><<<blocks, 256, 0, stream>>>(1.0, outer, sink);
  return cudaGetLastError();
}

}  // namespace
}  // namespace ub200

using namespace ub200;

extern "C" int usgscsm_b200_measure_fp64(int device, int op, double min_ms,
                                         double *instructions_per_second, double *sm_clock_mhz) {
  if (!instructions_per_second || op < 0 || op > 3 || !(min_ms >= 0.0))
    return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (device < 0 || device >= deviceCount())
    return fail(USGSCSM_B200_ENODEVICE, "no usable CUDA device with that index");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    UB_CUDA(cudaGetDeviceProperties(&prop, device));
    const int blocks = prop.multiProcessorCount * 4;
    double *sink = nullptr;
    UB_CUDA(cudaMalloc(&sink, sizeof(double)));
    cudaEvent_t e0, e1;
    UB_CUDA(cudaEventCreate(&e0));
    UB_CUDA(cudaEventCreate(&e1));
    auto launch = [&](int outer) -> cudaError_t {
      switch (op) {
        case 0: return runPeak<0>(blocks, outer, sink, nullptr);
        case 1: return runPeak<1>(blocks, outer, sink, nullptr);
        case 2: return runPeak<2>(blocks, outer, sink, nullptr);
        default: return runPeak<3>(blocks, outer, sink, nullptr);
      }
    };
    const int outer = 64;
    UB_CUDA(launch(outer));  // warm-up (clocks, instruction cache)
    UB_CUDA(cudaDeviceSynchronize());
    double totalMs = 0.0, instr = 0.0;
    const double perLaunch = (double)blocks * 256.0 * kChains * (double)kInner * outer;
    do {
      UB_CUDA(cudaEventRecord(e0));
      for (int r = 0; r < 8; r++) UB_CUDA(launch(outer));
      UB_CUDA(cudaEventRecord(e1));
      UB_CUDA(cudaEventSynchronize(e1));
      float ms = 0.f;
      UB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      totalMs += ms;
      instr += 8.0 * perLaunch;
    } while (totalMs < min_ms);
    *instructions_per_second = instr / (totalMs * 1e-3);
    if (sm_clock_mhz) *sm_clock_mhz = prop.clockRate / 1000.0;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    return USGSCSM_B200_OK;
  });
}
