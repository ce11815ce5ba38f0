// blocks_kernels.cu -- the building blocks of the per-point math as batch kernels:
// applyDistortion / removeDistortion (src/Distortion.cpp) and lagrangeInterp
// (src/Utilities.cpp:409-513), exported through the C-ABI so that the reference's own
// known-answer vectors (tests/DistortionTests.cpp, tests/UtilitiesTests.cpp) can be replayed
// against the device functions the sensor-model kernels are built from.
#include "kernels.h"
#include "device_math.cuh"

namespace ub200 {

namespace {

template <bool APPLY>
__global__ void distortionKernel(const __grid_constant__ FocalPlane fp, double precision, long n,
                                 const double *__restrict__ inX, const double *__restrict__ inY,
                                 double *__restrict__ outX, double *__restrict__ outY,
                                 uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double ox, oy;
    uint32_t st = 0;
    if (APPLY)
      applyDistortion(inX[i], inY[i], ox, oy, fp, precision, st);
    else
      removeDistortion(inX[i], inY[i], ox, oy, fp, precision, st);
    outX[i] = ox;
    outY[i] = oy;
    if (status) status[i] = st;
  }
}

template <int VLEN>
__global__ void lagrangeKernel(UniformTable tab, long n, const double *__restrict__ times,
                               double *__restrict__ out) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double v[VLEN];
    lagrangeInterp<VLEN>(tab, tab.values, times[i], v);
    for (int j = 0; j < VLEN; j++) out[i * VLEN + j] = v[j];
  }
}

int gridFor(long n) {
  long b = (n + 127) / 128;
  return (int)(b < 1 ? 1 : (b > 4096 ? 4096 : b));
}

}  // namespace

cudaError_t launchDistortion(bool apply, const FocalPlane &fp, double precision, long n,
                             const double *inX, const double *inY, double *outX, double *outY,
                             uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  if (apply)
    distortionKernel<true><<<gridFor(n), 128, 0, stream>>>(fp, precision, n, inX, inY, outX, outY, status);
  else
    distortionKernel<false><<<gridFor(n), 128, 0, stream>>>(fp, precision, n, inX, inY, outX, outY, status);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchLagrange(const UniformTable &tab, int vlen, long n, const double *times,
                           double *out, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  switch (vlen) {
    case 1: lagrangeKernel<1><<<gridFor(n), 128, 0, stream>>>(tab, n, times, out); break;
    case 2: lagrangeKernel<2><<<gridFor(n), 128, 0, stream>>>(tab, n, times, out); break;
    case 3: lagrangeKernel<3><<<gridFor(n), 128, 0, stream>>>(tab, n, times, out); break;
    case 4: lagrangeKernel<4><<<gridFor(n), 128, 0, stream>>>(tab, n, times, out); break;
    default: return cudaErrorInvalidValue;
  }
  countLaunch();
  return cudaGetLastError();
}

}  // namespace ub200
