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
__global__ void lagrangeKernel(UniformTable tab, LagrangeDivisors lag, long n, const double *__restrict__ times,
                               double *__restrict__ out) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double v[VLEN];
    lagrangeInterp<VLEN>(tab, lag, tab.values, times[i], v);
    for (int j = 0; j < VLEN; j++) out[i * VLEN + j] = v[j];
  }
}

// The focal-plane affine both ways, the look vector, the cubic and its roots: the free functions
// tests/UtilitiesTests.cpp:55-284, 443-475 exercise, over arrays.
__global__ void focalPlaneKernel(const __grid_constant__ FocalPlane fp, int toPixel, long n,
                                 const double *__restrict__ a, const double *__restrict__ b,
                                 double *__restrict__ outA, double *__restrict__ outB) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double u, v;
    if (toPixel)
      computePixel(a[i], b[i], fp, u, v);  // (focal x, y) -> (line, sample)
    else
      distortedFocalPlaneCoordinates(a[i], b[i], fp, u, v);  // (line, sample) -> (focal x, y)
    outA[i] = u;
    outB[i] = v;
  }
}

__global__ void lookVectorKernel(double zDirection, double focalLength, long n,
                                 const double *__restrict__ ux, const double *__restrict__ uy,
                                 double *__restrict__ look) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double l[3];
    cameraLookVector(ux[i], uy[i], zDirection, focalLength, l);
    look[3 * i] = l[0];
    look[3 * i + 1] = l[1];
    look[3 * i + 2] = l[2];
  }
}

struct Cubic {
  double c[4];
};

// mode 0: value / derivative at x; 1: polynomialRoot from guess a; 2: brentRoot between a and b
__global__ void cubicKernel(Cubic P, int mode, double eps, long n, const double *__restrict__ a,
                            const double *__restrict__ b, double *__restrict__ out0,
                            double *__restrict__ out1, uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    if (mode == 0) {
      if (out0) out0[i] = poly3(P.c, a[i]);
      if (out1) out1[i] = poly3Derivative(P.c, a[i]);
    } else if (mode == 1) {
      double root;
      const bool ok = polynomialRoot3(P.c, a[i], eps, root);
      out0[i] = ok ? root : nan("");
      if (status) status[i] = ok ? 0u : (uint32_t)USGSCSM_B200_ST_ERROR_UNKNOWN;
    } else {
      BrentSearch q;
      brentBegin(a[i], b[i], q);
      while (!brentStep([&](double t) { return poly3(P.c, t); }, eps, q)) {
      }
      out0[i] = q.failed ? nan("") : q.t;
      if (status)
        status[i] = q.status | (q.evals << USGSCSM_B200_ST_EVALS_SHIFT) |
                    (q.failed ? (uint32_t)USGSCSM_B200_ST_ERROR_UNKNOWN : 0u);
    }
  }
}

// AoS <-> SoA for the csm value types (csm::EcefCoord = 3 doubles, csm::ImageCoord = 2,
// csm/csm.h:212-308): arrays of structs are copied to the device as they are and split /
// joined there, so the per-point kernels always see coalesced SoA arrays.
template <int NC>
__global__ void aosSplitKernel(long n, const double *__restrict__ aos, double *__restrict__ c0,
                               double *__restrict__ c1, double *__restrict__ c2) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    c0[i] = aos[i * NC];
    c1[i] = aos[i * NC + 1];
    if (NC == 3) c2[i] = aos[i * NC + 2];
  }
}
template <int NC>
__global__ void aosJoinKernel(long n, const double *__restrict__ c0, const double *__restrict__ c1,
                              const double *__restrict__ c2, double *__restrict__ aos) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    aos[i * NC] = c0[i];
    aos[i * NC + 1] = c1[i];
    if (NC == 3) aos[i * NC + 2] = c2[i];
  }
}

// computeSensorPartials LS:1373-1396: (img1 - image_pt) / DELTA from the two searches.  The
// warning list of the scalar call only sees the adjusted search; a csm::Error thrown by
// either search escapes (NaN partials, error bit only).
__global__ void sensorPartialsCombineKernel(long n, const double *__restrict__ l0,
                                            const double *__restrict__ s0,
                                            const uint32_t *__restrict__ st0,
                                            const double *__restrict__ l1,
                                            const double *__restrict__ s1, double delta,
                                            double *__restrict__ dline, double *__restrict__ dsamp,
                                            double *__restrict__ achieved,
                                            uint32_t *__restrict__ status) {
  const uint32_t kErr = USGSCSM_B200_ST_ERROR_ALGORITHM | USGSCSM_B200_ST_ERROR_UNKNOWN |
                        USGSCSM_B200_ST_ERROR_OTHER;
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t a = st0 ? st0[i] : 0u, b = status[i];
    double pl = (l1[i] - l0[i]) / delta;
    double ps = (s1[i] - s0[i]) / delta;
    uint32_t st = (b & (USGSCSM_B200_ST_NO_INTERSECTION | USGSCSM_B200_ST_PRECISION_NOT_MET)) |
                  ((a | b) & kErr);
    if (st & kErr) {
      st &= kErr;
      pl = ps = nan("");
      if (achieved) achieved[i] = nan("");
    }
    if (!(isfinite(pl) && isfinite(ps))) st |= USGSCSM_B200_ST_NONFINITE;
    dline[i] = pl;
    dsamp[i] = ps;
    status[i] = st;
  }
}

// One parameter of computeAllSensorPartials (csm/RasterGM.cpp:101-140): the forward difference
// into slot k of the point's (dline, dsamp) list, achieved = max over the parameters starting
// from 0.0 ("prec > *achievedPrecision", so a NaN never wins), warnings accumulate.  On the
// last parameter a point with an error bit (the scalar call would have thrown) has its whole
// list set to NaN.
__global__ void sensorPartialsStepKernel(long n, const double *__restrict__ l0,
                                         const double *__restrict__ s0,
                                         const uint32_t *__restrict__ st0,
                                         const double *__restrict__ l1,
                                         const double *__restrict__ s1,
                                         const double *__restrict__ achP,
                                         const uint32_t *__restrict__ stP, double delta,
                                         double *__restrict__ partials, int numParams, int k,
                                         double *__restrict__ achieved,
                                         uint32_t *__restrict__ status) {
  const uint32_t kErr = USGSCSM_B200_ST_ERROR_ALGORITHM | USGSCSM_B200_ST_ERROR_UNKNOWN |
                        USGSCSM_B200_ST_ERROR_OTHER;
  const uint32_t kKeep = kErr | USGSCSM_B200_ST_NO_INTERSECTION | USGSCSM_B200_ST_PRECISION_NOT_MET;
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t acc = k == 0 ? (st0 ? st0[i] & kKeep : 0u) : status[i];
    double am = (k == 0 || !achieved) ? 0.0 : achieved[i];
    acc |= stP[i] & kKeep;
    const double a = achP[i];
    if (a > am) am = a;
    const double pl = (l1[i] - l0[i]) / delta;
    const double ps = (s1[i] - s0[i]) / delta;
    double *out = partials + (size_t)i * numParams * 2;
    out[2 * k] = pl;
    out[2 * k + 1] = ps;
    if (!(isfinite(pl) && isfinite(ps))) acc |= USGSCSM_B200_ST_NONFINITE;
    if (k == numParams - 1 && (acc & kErr)) {
      acc = (acc & kErr) | USGSCSM_B200_ST_NONFINITE;
      for (int q = 0; q < 2 * numParams; q++) out[q] = nan("");
      am = nan("");
    }
    status[i] = acc;
    if (achieved) achieved[i] = am;
  }
}

// r[i] = refinedReciprocal(b[i]); a divisor the fast path can never accept (zero, subnormal,
// huge, non-finite) gets NaN, which sends every division by it to the slow path.
__global__ void reciprocalKernel(int n, const double *__restrict__ b, double *__restrict__ r) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) r[i] = refinedReciprocal(b[i]);
}

// divideBy(a, b, refinedReciprocal(b)) against a / b on pseudo-random operands over the
// whole exponent range (every 16th pair has extreme exponents).  Counts bit mismatches.
__device__ inline unsigned long long mix64(unsigned long long z) {
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
__global__ void divisionSelfTestKernel(unsigned long long seed, long perThread,
                                       unsigned long long *mismatches) {
  unsigned long long s = seed + ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * 0x632be59bd9b4e019ull;
  unsigned long long bad = 0;
  const double fixed[8] = {5040.0, 720.0, 240.0, 144.0, 120.0, 24.0, 12.0, 6.0};
  for (long k = 0; k < perThread; k++) {
    unsigned long long ra = mix64(s += 1), rb = mix64(s += 1);
    double a, b;
    if ((k & 15) == 0) {  // raw bit patterns: every exponent, subnormals, inf, NaN
      a = __longlong_as_double((long long)ra);
      b = __longlong_as_double((long long)rb);
    } else {  // moderate exponents (+-2^40), full random mantissas
      a = __longlong_as_double((long long)((ra & 0x800fffffffffffffull) | ((0x3d7ull + (ra >> 52) % 80) << 52)));
      b = __longlong_as_double((long long)((rb & 0x800fffffffffffffull) | ((0x3d7ull + (rb >> 52) % 80) << 52)));
      if ((k & 15) == 1) b = fixed[rb & 7];
    }
    double want = a / b;
    double got = divideBy(a, b, refinedReciprocal(b));
    if (__double_as_longlong(want) != __double_as_longlong(got) && !(want != want && got != got)) bad++;
  }
  if (bad) atomicAdd(mismatches, bad);
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

cudaError_t launchLagrange(const UniformTable &tab, const LagrangeDivisors &lag, int vlen, long n, const double *times,
                           double *out, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  switch (vlen) {
    case 1: lagrangeKernel<1><<<gridFor(n), 128, 0, stream>>>(tab, lag, n, times, out); break;
    case 2: lagrangeKernel<2><<<gridFor(n), 128, 0, stream>>>(tab, lag, n, times, out); break;
    case 3: lagrangeKernel<3><<<gridFor(n), 128, 0, stream>>>(tab, lag, n, times, out); break;
    case 4: lagrangeKernel<4><<<gridFor(n), 128, 0, stream>>>(tab, lag, n, times, out); break;
    default: return cudaErrorInvalidValue;
  }
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchReciprocals(int n, const double *b, double *r, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  reciprocalKernel<<<(n + 63) / 64, 64, 0, stream>>>(n, b, r);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchDivisionSelfTest(unsigned long long seed, int blocks, long perThread,
                                   unsigned long long *mismatches, cudaStream_t stream) {
  divisionSelfTestKernel<<<blocks, 256, 0, stream>>>(seed, perThread, mismatches);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchAosSplit(int ncomp, long n, const double *aos, double *c0, double *c1,
                           double *c2, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  if (ncomp == 3)
    aosSplitKernel<3><<<blocks, 256, 0, stream>>>(n, aos, c0, c1, c2);
  else
    aosSplitKernel<2><<<blocks, 256, 0, stream>>>(n, aos, c0, c1, c2);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchAosJoin(int ncomp, long n, const double *c0, const double *c1, const double *c2,
                          double *aos, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  if (ncomp == 3)
    aosJoinKernel<3><<<blocks, 256, 0, stream>>>(n, c0, c1, c2, aos);
  else
    aosJoinKernel<2><<<blocks, 256, 0, stream>>>(n, c0, c1, c2, aos);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchSensorPartialsCombine(long n, const double *l0, const double *s0,
                                        const uint32_t *st0, const double *l1, const double *s1,
                                        double delta, double *dline, double *dsamp,
                                        double *achieved, uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  sensorPartialsCombineKernel<<<blocks, 256, 0, stream>>>(n, l0, s0, st0, l1, s1, delta, dline,
                                                          dsamp, achieved, status);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchSensorPartialsStep(long n, const double *l0, const double *s0,
                                     const uint32_t *st0, const double *l1, const double *s1,
                                     const double *achP, const uint32_t *stP, double delta,
                                     double *partials, int numParams, int k, double *achieved,
                                     uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  sensorPartialsStepKernel<<<blocks, 256, 0, stream>>>(n, l0, s0, st0, l1, s1, achP, stP, delta,
                                                       partials, numParams, k, achieved, status);
  countLaunch();
  return cudaGetLastError();
}

static int katBlocks(long n) { return (int)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184); }

cudaError_t launchFocalPlane(const FocalPlane &fp, int toPixel, long n, const double *a, const double *b,
                             double *outA, double *outB, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  focalPlaneKernel<<<katBlocks(n), 256, 0, stream>>>(fp, toPixel, n, a, b, outA, outB);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchLookVector(double zDirection, double focalLength, long n, const double *ux,
                             const double *uy, double *look, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  lookVectorKernel<<<katBlocks(n), 256, 0, stream>>>(zDirection, focalLength, n, ux, uy, look);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchCubic(const double coeffs[4], int mode, double eps, long n, const double *a,
                        const double *b, double *out0, double *out1, uint32_t *status,
                        cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  Cubic P;
  for (int k = 0; k < 4; k++) P.c[k] = coeffs[k];
  cubicKernel<<<katBlocks(n), 256, 0, stream>>>(P, mode, eps, n, a, b, out0, out1, status);
  countLaunch();
  return cudaGetLastError();
}

}  // namespace ub200
