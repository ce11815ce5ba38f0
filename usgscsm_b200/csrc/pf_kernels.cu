// pf_kernels.cu -- push-frame groundToImage / imageToGround, one thread per point, FP64.
//
// Replaces UsgsAstroPushFrameSensorModel::groundToImage
// (src/UsgsAstroPushFrameSensorModel.cpp:715-778, with calcFrameDistance :2496-2521) and
// ::imageToGround (:842-870, losToEcf :1726-1824, losEllipsoidIntersect :1883-1950).
//
// B200-first restructuring (bit-identical to the per-point code): a push-frame image has
// ONE exposure time per framelet (getImageTime, PF:1258-1276), so everything the reference
// recomputes per point and per secant step from the time alone -- two Lagrange-8
// interpolations of the ephemeris, the in-track / cross-track / radial frame, the
// Lagrange-8 quaternion, its rotation matrix and the attitude correction -- takes at most
// numFramelets + 2 distinct values.  A set-up kernel evaluates them once per model with the
// same device functions (same arithmetic, same rounding => same bits); the per-point
// kernels read the 12 (or 21) doubles of their framelet from a copy staged in shared
// memory.  A framelet number outside the table (imageToGround of a line far outside the
// image) falls back to the in-thread evaluation.
#include "kernels.h"
#include "ls_common.cuh"
#include "work_queue.cuh"

// resident blocks of 256 threads per SM the register allocation is held to (measured best)
#ifndef UB_PF_G2I_MINB
#define UB_PF_G2I_MINB 4
#endif
#ifndef UB_PF_I2G_MINB
#define UB_PF_I2G_MINB 3
#endif

namespace ub200 {

namespace {

constexpr int kThreads = 256;

// The time framelet number of getImageTime (PF:1258-1276) for an image line.
UB_DEV int pfTimeFramelet(const LsState &S, double line) {
  int frameletNumber = truncToInt(line / S.summedFrameletHeight);
  if (S.frameletsOrderReversed) frameletNumber = S.numFramelets - frameletNumber - 1;
  return frameletNumber;
}

UB_DEV double pfFrameletTime(const LsState &S, int frameletNumber) {
  double time = S.startingEphemerisTime + 0.5 * S.exposureDuration +
                frameletNumber * S.interframeDelay;
  time -= S.centerEphemerisTime;
  return time;
}

struct Orientation {
  double pos[3];
  double camToBody[9];
  double attCorr[9];  // valid only when S.hasAttAdj
};

// getAdjSensorPosVel PF:1955-2034 + getQuaternions PF:1679-1692 +
// calculateRotationMatrixFromQuaternions + calculateAttitudeCorrection at one time.
UB_DEV void computeOrientation(const LsState &S, const OrientationTables &T, double time,
                               Orientation &o) {
  double vel[3];
  adjSensorPosVel<true>(S, T, time, o.pos, vel);
  cameraToBodyMatrix(S, T, time, o.camToBody);
  if (S.hasAttAdj) attitudeCorrection(S, time, o.attCorr);
}

// Where the per-framelet orientations are read from.
struct OrientCache {
  const double *entries;  // entry k + 1 <-> time framelet number k; null = no cache
  int stride;             // doubles per entry in `entries`
  bool pairs;             // entries of 12 doubles on 16-byte boundaries (the staged copy): 128-bit loads
};

// ATT: -1 = S.hasAttAdj decides at run time, 0 / 1 = the launcher knows (compile-time variant).
template <int ATT>
UB_DEV bool pfHasAttAdj(const LsState &S) { return ATT < 0 ? S.hasAttAdj != 0 : ATT != 0; }

// entry of a framelet the caller knows to be inside the table
template <int ATT>
UB_DEV void orientationFromTable(const LsState &S, const OrientCache &C, int frameletNumber,
                                 Orientation &o) {
  if (ATT == 0 && C.pairs) {
    const double2 *q = reinterpret_cast<const double2 *>(C.entries + (long)(frameletNumber + 1) * 12);
    const double2 a = q[0], b = q[1], c = q[2], d = q[3], f = q[4], g = q[5];
    o.pos[0] = a.x, o.pos[1] = a.y, o.pos[2] = b.x;
    o.camToBody[0] = b.y, o.camToBody[1] = c.x, o.camToBody[2] = c.y;
    o.camToBody[3] = d.x, o.camToBody[4] = d.y, o.camToBody[5] = f.x;
    o.camToBody[6] = f.y, o.camToBody[7] = g.x, o.camToBody[8] = g.y;
    return;
  }
  const double *e = C.entries + (long)(frameletNumber + 1) * C.stride;
#pragma unroll
  for (int i = 0; i < 3; i++) o.pos[i] = e[i];
#pragma unroll
  for (int i = 0; i < 9; i++) o.camToBody[i] = e[3 + i];
  if (pfHasAttAdj<ATT>(S)) {
#pragma unroll
    for (int i = 0; i < 9; i++) o.attCorr[i] = e[12 + i];
  }
}

template <int ATT = -1>
UB_DEV void orientationAt(const LsState &S, const OrientationTables &T, const OrientCache &C,
                          int frameletNumber, Orientation &o) {
  if (C.entries && frameletNumber >= -1 && frameletNumber <= S.numFramelets) {
    orientationFromTable<ATT>(S, C, frameletNumber, o);
  } else {
    computeOrientation(S, T, pfFrameletTime(S, frameletNumber), o);
  }
}

// Second half of computeDetectorView PF:2038-2110 given the exterior orientation.
template <int ATT = -1>
UB_DEV void viewFromOrientation(const LsState &S, const Orientation &o, double gx, double gy,
                                double gz, double &focalX, double &focalY) {
  double bodyLook[3] = {gx - o.pos[0], gy - o.pos[1], gz - o.pos[2]};
  double cameraLook[3];
  matTVec(o.camToBody, bodyLook, cameraLook);
  double adjustedLook[3];
  if (pfHasAttAdj<ATT>(S)) {
    matTVec(o.attCorr, cameraLook, adjustedLook);
  } else {  // Euler angles identically 0: the correction is the identity
    adjustedLook[0] = cameraLook[0];
    adjustedLook[1] = cameraLook[1];
    adjustedLook[2] = cameraLook[2];
  }
  double lookScale = (S.fp.focalLength + param(S, 15)) / adjustedLook[2];
  focalX = adjustedLook[0] * lookScale;
  focalY = adjustedLook[1] * lookScale;
}

// Stage the framelet orientations into shared memory (12 doubles per entry, or all 21 when
// the attitude correction is live).  All threads of the block call it.
__device__ inline OrientCache stageOrientCache(const LsState &S, double *smem, bool stage) {
  OrientCache c;
  c.pairs = false;
  if (!S.frameletOrient) {
    c.entries = nullptr;
    c.stride = 0;
    return c;
  }
  if (!stage) {
    c.entries = S.frameletOrient;
    c.stride = kFrameletOrientDoubles;
    return c;
  }
  const int stride = S.hasAttAdj ? kFrameletOrientDoubles : 12;
  const int entries = S.numFramelets + 2;
  for (int i = threadIdx.x; i < entries * stride; i += blockDim.x) {
    int e = i / stride, j = i - e * stride;
    smem[i] = __ldg(S.frameletOrient + (long)e * kFrameletOrientDoubles + j);
  }
  __syncthreads();
  c.entries = smem;
  c.stride = stride;
  c.pairs = stride == 12;
  return c;
}

__host__ __device__ inline size_t orientCacheBytes(const LsState &S) {
  return (size_t)(S.numFramelets + 2) * (S.hasAttAdj ? kFrameletOrientDoubles : 12) *
         sizeof(double);
}

__global__ void pfFrameletOrientKernel(const __grid_constant__ LsState S, double *out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;  // entry index; time framelet number k - 1
  if (k >= S.numFramelets + 2) return;
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  Orientation o;
#pragma unroll
  for (int i = 0; i < 9; i++) o.attCorr[i] = 0.0;
  computeOrientation(S, T, pfFrameletTime(S, k - 1), o);
  double *e = out + (long)k * kFrameletOrientDoubles;
  for (int i = 0; i < 3; i++) e[i] = o.pos[i];
  for (int i = 0; i < 9; i++) e[3 + i] = o.camToBody[i];
  for (int i = 0; i < 9; i++) e[12 + i] = o.attCorr[i];
}

// ---- groundToImage -------------------------------------------------------------------------
// DIST >= 0 / ATT >= 0: compile-time variants of the launcher (distortion type, attitude
// correction live or not).  The launcher picks them only when the framelet table exists and
// (2 numFramelets + 1) * summedFrameletHeight < 2^53 (pfFrameletNumbersExact), and then the time
// framelet of calcFrameDistance needs no division and is always inside the table:
// (e + 0.5) * h is exact, so its correctly rounded quotient by h is e + 0.5 and truncates to e,
// and (h * e + 0.5) / h = e + 0.5 / h truncates to e as well -- the final view is the last
// evaluation's.  e is clamped to [0, numFramelets], the table spans [-1, numFramelets].
template <int DIST = -1, int ATT = -1>
UB_DEV void pfGroundToImagePoint(const LsState &S, const OrientationTables &T,
                                 const OrientCache &C, double gx, double gy, double gz,
                                 double desiredPrecision, double &outLine, double &outSamp,
                                 double &achieved, uint32_t &status) {
  const FocalPlane &fp = S.fp;
  const int summedFrameletHeight = S.summedFrameletHeight;
  const int numFramelets = S.numFramelets;

  int startFramelet = 0;
  double startDistance = 0.0;
  int endFramelet = 0;
  double endDistance = 0.0;
  double focalX = 0.0, focalY = 0.0;
  int viewFramelet = 0;  // time framelet number focalX/focalY were evaluated at
  int count = 0;
  // Evaluation order of the reference: calcFrameDistance(0), calcFrameDistance(n - 1), then
  // one per secant step.  stage 0/1 are the two seeds.
  int stage = 0;
  uint32_t evals = 0;
  for (;;) {
    evals++;
    // calcFrameDistance PF:2496-2521 at endFramelet
    Orientation o;
    if (DIST >= 0) {
      viewFramelet = S.frameletsOrderReversed ? numFramelets - endFramelet - 1 : endFramelet;
      orientationFromTable<ATT>(S, C, viewFramelet, o);
    } else {
      double centerLine = (endFramelet + 0.5) * summedFrameletHeight;
      viewFramelet = pfTimeFramelet(S, centerLine);
      orientationAt<ATT>(S, T, C, viewFramelet, o);
    }
    viewFromOrientation<ATT>(S, o, gx, gy, gz, focalX, focalY);
    double detectorLine = fp.iTransL[0] + fp.iTransL[1] * focalX + fp.iTransL[2] * focalY +
                          fp.detectorLineOrigin - fp.startingDetectorLine;
    detectorLine = divideBy(detectorLine, fp.lineSumming);
    detectorLine -= S.halfOverlap;
    double distance = detectorLine - summedFrameletHeight / 2.0;
    if (stage == 0) {
      startFramelet = 0;
      startDistance = distance;
      endFramelet = numFramelets - 1;
      stage = 1;
      continue;
    }
    endDistance = distance;
    // while (startFramelet != endFramelet && startDistance != endDistance && count++ < 20)
    if (!(startFramelet != endFramelet)) break;
    if (!(startDistance != endDistance)) {
      status |= USGSCSM_B200_ST_SEARCH_STALL;
      break;
    }
    if (!(count++ < 20)) {
      status |= USGSCSM_B200_ST_ITERATION_CAP;
      break;
    }
    double offset = endDistance * (endFramelet - startFramelet) / (endDistance - startDistance);
    startFramelet = endFramelet;
    startDistance = endDistance;
    endFramelet = truncToInt(round(endFramelet - offset));
    endFramelet = max(0, min(numFramelets, endFramelet));
  }

  // getImageTime(ImageCoord(0.5 + summedFrameletHeight * endFramelet, 0.0)) + computeDetectorView:
  // the same framelet time as the last distance evaluation, hence the same focal-plane point.
  status |= evals << USGSCSM_B200_ST_EVALS_SHIFT;
  if (DIST < 0) {
    int finalFramelet = pfTimeFramelet(S, 0.5 + summedFrameletHeight * endFramelet);
    if (finalFramelet != viewFramelet) {
      Orientation o;
      orientationAt<ATT>(S, T, C, finalFramelet, o);
      viewFromOrientation<ATT>(S, o, gx, gy, gz, focalX, focalY);
    }
  }
  double distortedFocalX, distortedFocalY;
  applyDistortion<DIST>(focalX, focalY, distortedFocalX, distortedFocalY, fp, desiredPrecision, status);
  double detectorLine =
      fp.iTransL[0] + fp.iTransL[1] * distortedFocalX + fp.iTransL[2] * distortedFocalY;
  double detectorSample =
      fp.iTransS[0] + fp.iTransS[1] * distortedFocalX + fp.iTransS[2] * distortedFocalY;
  double line =
      divideBy(detectorLine + fp.detectorLineOrigin - fp.startingDetectorLine, fp.lineSumming);
  line += endFramelet * summedFrameletHeight;
  line -= S.halfOverlap;
  outLine = line;
  outSamp = divideBy(detectorSample + fp.detectorSampleOrigin - fp.startingDetectorSample,
                     fp.sampleSumming);
  achieved = desiredPrecision;  // PF:769-771
}

UB_DEV uint32_t pfImageRangeStatus(double line, double samp, const LsState &S) {
  uint32_t st = 0;
  if (!(isfinite(line) && isfinite(samp))) st |= USGSCSM_B200_ST_NONFINITE;
  // getValidImageRange PF:1465-1472: (0,0) .. (nReducedLines, nSamples)
  if (!(line >= 0.0 && line <= (double)S.nReducedLines && samp >= 0.0 &&
        samp <= (double)S.nSamples))
    st |= USGSCSM_B200_ST_OUT_OF_IMAGE;
  return st;
}

template <bool STAGE, int DIST, int ATT>
__global__ void __launch_bounds__(kThreads, UB_PF_G2I_MINB)
    pfGroundToImageKernel(const __grid_constant__ LsState S, long n, const double *__restrict__ x,
                          const double *__restrict__ y, const double *__restrict__ z, double prec,
                          double *__restrict__ line, double *__restrict__ samp,
                          double *__restrict__ achieved, uint32_t *__restrict__ status,
                          unsigned long long *__restrict__ next, int grab) {
  extern __shared__ __align__(16) double smem[];
  OrientCache C = stageOrientCache(S, smem, STAGE);
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  forEachPointDynamic(n, next, grab, [&](long i) {
    double gx = __ldcs(x + i), gy = __ldcs(y + i), gz = __ldcs(z + i);
    double l, s, a;
    uint32_t st = 0;
    pfGroundToImagePoint<DIST, ATT>(S, T, C, gx, gy, gz, prec, l, s, a, st);
    st |= pfImageRangeStatus(l, s, S);
    __stcs(line + i, l);
    __stcs(samp + i, s);
    if (achieved) __stcs(achieved + i, a);
    if (status) __stcs(status + i, st);
  });
}

// ---- imageToGround -------------------------------------------------------------------------
// A framelet outside the table (a line far outside the image): evaluated in the thread, out of
// line so that the specialised kernels do not carry its registers and instructions.
static __device__ __noinline__ void computeOrientationCold(const LsState &S, const OrientationTables &T,
                                                           int frameletNumber, Orientation &o) {
  computeOrientation(S, T, pfFrameletTime(S, frameletNumber), o);
}

// line / h and fmod(line, h) for the framelet height h (a positive integer as a double) from ONE
// quotient with the reciprocal rh = refinedReciprocal(h) hoisted: q = RN(|line| / h) (divideBy
// returns the division's bits) truncates to n or n + 1, n the integer part of the true quotient;
// |line| - trunc(q) h is exact in one fma (a multiple of ulp(line) below 2 h), and one
// correction brings it into [0, h): that is fmod's exact remainder, with line's sign.
UB_DEV void pfSplitLine(double line, double h, double rh, double &quotient, double &remainder) {
  const double a = fabs(line);
  if (!(a < 4503599627370496.0)) {  // 2^52 and up, NaN: the library routines
    quotient = line / h;
    remainder = fmod(line, h);
    return;
  }
  const double q = divideBy(a, h, rh);
  double r = __fma_rn(-trunc(q), h, a);
  if (r < 0.0) r += h;
  quotient = copysign(q, line);
  remainder = copysign(r, line);
}

// DIST >= 0 / ATT >= 0: the launcher's compile-time variants (see pfGroundToImagePoint); rh is
// refinedReciprocal(summedFrameletHeight), used by them only.
template <int DIST = -1, int ATT = -1>
UB_DEV void pfImageToGroundPoint(const LsState &S, const OrientationTables &T,
                                 const OrientCache &C, double line, double sample, double height,
                                 double desiredPrecision, double &x, double &y, double &z,
                                 double &achieved, uint32_t &status, double rh = 0.0) {
  int frameletNumber;
  double frameletLine;  // line within the framelet, PF:1757-1764
  Orientation o;
  if (DIST >= 0) {
    double quotient;
    pfSplitLine(line, (double)S.summedFrameletHeight, rh, quotient, frameletLine);
    frameletNumber = truncToInt(quotient);
    if (S.frameletsOrderReversed) frameletNumber = S.numFramelets - frameletNumber - 1;
    if (C.entries && frameletNumber >= -1 && frameletNumber <= S.numFramelets) {
      orientationFromTable<ATT>(S, C, frameletNumber, o);
    } else {
      Orientation cold;  // the callee's copy lives in local memory; `o` stays in registers
      computeOrientationCold(S, T, frameletNumber, cold);
#pragma unroll
      for (int i = 0; i < 3; i++) o.pos[i] = cold.pos[i];
#pragma unroll
      for (int i = 0; i < 9; i++) o.camToBody[i] = cold.camToBody[i];
#pragma unroll
      for (int i = 0; i < 9; i++) o.attCorr[i] = cold.attCorr[i];
    }
  } else {
    frameletNumber = pfTimeFramelet(S, line);
    orientationAt<ATT>(S, T, C, frameletNumber, o);
    frameletLine = fmod(line, (double)S.summedFrameletHeight);
  }
  if (S.frameletsFlipped) frameletLine = S.summedFrameletHeight - frameletLine;
  frameletLine += S.halfOverlap;

  double distortedFocalPlaneX, distortedFocalPlaneY;
  distortedFocalPlaneCoordinates(frameletLine, sample, S.fp, distortedFocalPlaneX,
                                 distortedFocalPlaneY);
  double undistortedFocalPlaneX, undistortedFocalPlaneY;
  removeDistortion<DIST>(distortedFocalPlaneX, distortedFocalPlaneY, undistortedFocalPlaneX,
                         undistortedFocalPlaneY, S.fp, 1.0E-6, status);
  double cameraLook[3];
  cameraLookVector(undistortedFocalPlaneX, undistortedFocalPlaneY, S.fp.zDirection,
                   S.fp.focalLength * (1 - param(S, 15) / S.halfSwath), cameraLook);
  double corrected[3];
  if (pfHasAttAdj<ATT>(S)) {
    matVec(o.attCorr, cameraLook, corrected);
  } else {
    corrected[0] = cameraLook[0];
    corrected[1] = cameraLook[1];
    corrected[2] = cameraLook[2];
  }
  double bodyLook[3];
  matVec(o.camToBody, corrected, bodyLook);
  lsEllipsoidIntersect(S.ell, height, o.pos, bodyLook, desiredPrecision, x, y, z, achieved,
                       status);
  if ((desiredPrecision > 0.0) && (achieved > desiredPrecision))
    status |= USGSCSM_B200_ST_PRECISION_NOT_MET;
  if (status & USGSCSM_B200_ST_ERROR_ALGORITHM) {
    status = USGSCSM_B200_ST_ERROR_ALGORITHM;
    x = y = z = achieved = nan("");
  }
  if (!finite3(x, y, z)) status |= USGSCSM_B200_ST_NONFINITE;
}

template <bool STAGE, bool GRID, int DIST, int ATT>
__global__ void __launch_bounds__(kThreads, UB_PF_I2G_MINB)
    pfImageToGroundKernel(const __grid_constant__ LsState S, long n, const double *__restrict__ line,
                          const double *__restrict__ samp, GridSpec g,
                          const double *__restrict__ height, int heightStride, double prec,
                          double *__restrict__ x, double *__restrict__ y, double *__restrict__ z,
                          double *__restrict__ achieved, uint32_t *__restrict__ status) {
  extern __shared__ __align__(16) double smem[];
  OrientCache C = stageOrientCache(S, smem, STAGE);
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  const double rh = DIST >= 0 ? refinedReciprocal((double)S.summedFrameletHeight) : 0.0;
  const long stride = (long)gridDim.x * blockDim.x;
  const long first = (long)blockIdx.x * blockDim.x + threadIdx.x;
  // dense grid: (row, column) of the thread's point, advanced by the stride's (rows, columns)
  // instead of a 64-bit division per point
  long r = 0, c = 0, strideRows = 0, strideCols = 0;
  if (GRID) {
    r = first / g.nSamps;
    c = first - r * g.nSamps;
    strideRows = stride / g.nSamps;
    strideCols = stride - strideRows * g.nSamps;
  }
  for (long i = first; i < n; i += stride) {
    double l, s;
    if (GRID) {
      l = g.line0 + (double)(g.row0 + r) * g.dline;
      s = g.samp0 + (double)c * g.dsamp;
      r += strideRows;
      c += strideCols;
      if (c >= g.nSamps) {
        c -= g.nSamps;
        r++;
      }
    } else {
      l = __ldcs(line + i);
      s = __ldcs(samp + i);
    }
    double h = heightStride ? __ldcs(height + i) : __ldg(height);
    double px, py, pz, a;
    uint32_t st = 0;
    pfImageToGroundPoint<DIST, ATT>(S, T, C, l, s, h, prec, px, py, pz, a, st, rh);
    __stcs(x + i, px);
    __stcs(y + i, py);
    __stcs(z + i, pz);
    if (achieved) __stcs(achieved + i, a);
    if (status) __stcs(status + i, st);
  }
}

// computeGroundPartials PF:1091-1116: as the line-scan model's, delta = m_gsd.
template <bool STAGE>
__global__ void __launch_bounds__(kThreads)
    pfGroundPartialsKernel(const __grid_constant__ LsState S, long n, const double *__restrict__ x,
                           const double *__restrict__ y, const double *__restrict__ z,
                           double *__restrict__ partials, uint32_t *__restrict__ status,
                           unsigned long long *__restrict__ next, int grab) {
  extern __shared__ __align__(16) double smem[];
  OrientCache C = stageOrientCache(S, smem, STAGE);
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  forEachPointDynamic(n, next, grab, [&](long i) {
    const double g[3] = {__ldcs(x + i), __ldcs(y + i), __ldcs(z + i)};
    const double delta = S.gsd;
    double l[4], s[4];
    uint32_t st = 0;
#pragma unroll 1
    for (int k = 0; k < 4; k++) {
      double q[3] = {g[0], g[1], g[2]};
      if (k) q[k - 1] = g[k - 1] + delta;
      double a;
      uint32_t sk = 0;
      pfGroundToImagePoint(S, T, C, q[0], q[1], q[2], 0.001, l[k], s[k], a, sk);
      st |= sk;
    }
    double p[6];
    p[0] = (l[1] - l[0]) / delta;
    p[3] = (s[1] - s[0]) / delta;
    p[1] = (l[2] - l[0]) / delta;
    p[4] = (s[2] - s[0]) / delta;
    p[2] = (l[3] - l[0]) / delta;
    p[5] = (s[3] - s[0]) / delta;
    storePartials(partials, status, i, p, st);
  });
}

// imageToRemoteImagingLocus PF:1060-1082: losToEcf (PF:1726-1824) at the pixel; point = sensor
// position of the pixel's framelet, direction = minus the body-fixed look vector.
UB_DEV void pfRemoteLocusPoint(const LsState &S, const OrientationTables &T, const OrientCache &C,
                               double line, double sample, double pos[3], double dir[3],
                               uint32_t &status) {
  int frameletNumber = pfTimeFramelet(S, line);
  Orientation o;
  orientationAt(S, T, C, frameletNumber, o);
  double frameletLine = fmod(line, (double)S.summedFrameletHeight);
  if (S.frameletsFlipped) frameletLine = S.summedFrameletHeight - frameletLine;
  frameletLine += S.halfOverlap;
  double distortedFocalPlaneX, distortedFocalPlaneY;
  distortedFocalPlaneCoordinates(frameletLine, sample, S.fp, distortedFocalPlaneX,
                                 distortedFocalPlaneY);
  double undistortedFocalPlaneX, undistortedFocalPlaneY;
  removeDistortion(distortedFocalPlaneX, distortedFocalPlaneY, undistortedFocalPlaneX,
                   undistortedFocalPlaneY, S.fp, 1.0E-6, status);
  double cameraLook[3];
  cameraLookVector(undistortedFocalPlaneX, undistortedFocalPlaneY, S.fp.zDirection,
                   S.fp.focalLength * (1 - param(S, 15) / S.halfSwath), cameraLook);
  double corrected[3];
  if (S.hasAttAdj) {
    matVec(o.attCorr, cameraLook, corrected);
  } else {
    corrected[0] = cameraLook[0];
    corrected[1] = cameraLook[1];
    corrected[2] = cameraLook[2];
  }
  double bodyLook[3];
  matVec(o.camToBody, corrected, bodyLook);
  for (int k = 0; k < 3; k++) {
    pos[k] = o.pos[k];
    dir[k] = -bodyLook[k];
  }
  if (status & USGSCSM_B200_ST_ERROR_ALGORITHM) {
    status = USGSCSM_B200_ST_ERROR_ALGORITHM;
    pos[0] = pos[1] = pos[2] = dir[0] = dir[1] = dir[2] = nan("");
  }
  if (!(finite3(pos[0], pos[1], pos[2]) && finite3(dir[0], dir[1], dir[2])))
    status |= USGSCSM_B200_ST_NONFINITE;
}

template <bool STAGE>
__global__ void __launch_bounds__(kThreads)
    pfRemoteLocusKernel(const __grid_constant__ LsState S, long n, const double *__restrict__ line,
                        const double *__restrict__ samp, LocusOut o) {
  extern __shared__ __align__(16) double smem[];
  OrientCache C = stageOrientCache(S, smem, STAGE);
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double pos[3], dir[3];
    uint32_t st = 0;
    pfRemoteLocusPoint(S, T, C, __ldcs(line + i), __ldcs(samp + i), pos, dir, st);
    storeLocus(o, i, pos, dir, 0.0, st);
  }
}

// imageToProximateImagingLocus PF:995-1055 (same construction as the line-scan model's).
template <bool STAGE>
__global__ void __launch_bounds__(kThreads)
    pfProximateLocusKernel(const __grid_constant__ LsState S, long n, const double *__restrict__ line,
                           const double *__restrict__ samp, const double *__restrict__ x,
                           const double *__restrict__ y, const double *__restrict__ z, double prec,
                           LocusOut o) {
  extern __shared__ __align__(16) double smem[];
  OrientCache C = stageOrientCache(S, smem, STAGE);
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double l = __ldcs(line + i), sm = __ldcs(samp + i);
    double pos[3], dir[3], a;
    uint32_t st = 0;
    proximateLocusFromImageToGround(
        S.ell, S.gsd, __ldcs(x + i), __ldcs(y + i), __ldcs(z + i), prec,
        [&](double h, double g[3], double &ach, uint32_t &s) {
          pfImageToGroundPoint(S, T, C, l, sm, h, prec, g[0], g[1], g[2], ach, s);
        },
        pos, dir, a, st);
    storeLocus(o, i, pos, dir, a, st);
  }
}

// getSensorPosition(double time) PF:1290-1303: getAdjSensorPosVel at `time`.
__global__ void pfSensorPositionKernel(const __grid_constant__ LsState S, double time,
                                       double *__restrict__ xyz) {
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  double pos[3], vel[3];
  adjSensorPosVel<true>(S, T, time, pos, vel);
  xyz[0] = pos[0];
  xyz[1] = pos[1];
  xyz[2] = pos[2];
}

// getImageTime PF:1258-1276, getSensorPosition(ImageCoord) PF:1281-1299,
// getSensorVelocity(ImageCoord) PF:1303-1320, per image point.
__global__ void pfSensorStateKernel(const __grid_constant__ LsState S, long n,
                                    const double *__restrict__ line, SensorStateOut o) {
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double time = o.atTime ? line[i] : pfFrameletTime(S, pfTimeFramelet(S, line[i]));
    double pos[3], vel[3];
    adjSensorPosVel<true>(S, T, time, pos, vel);
    storeSensorState(o, i, time, pos, vel, 0u);
  }
}

LaunchShape pfShape(const LsState &S, long n, int smCount, size_t &smemBytes, bool &stage,
                    bool staticSplit = false) {
  size_t bytes = S.frameletOrient ? orientCacheBytes(S) : 0;
  stage = bytes > 0 && bytes <= kMaxStagedTableBytes;
  smemBytes = stage ? bytes : 0;
  LaunchShape sh;
  sh.threads = kThreads;
  long blocksNeeded = (n + kThreads - 1) / kThreads;
  // staticSplit: the kernel splits its points statically over the grid (imageToGround); two
  // short waves of blocks (USGSCSM_B200_PF_WAVES) balance the tail of a sub-millisecond launch
  // (one 4096^2 image: +10 %).  The searches take their points from a queue: one wave.
  static const int envWaves = [] {
    const char *v = getenv("USGSCSM_B200_PF_WAVES");
    const int w = v && *v ? atoi(v) : 2;
    return w < 1 ? 1 : w;
  }();
  const int waves = staticSplit ? envWaves : 1;
  long persistent = (long)smCount * (UB_PF_G2I_MINB > kBlocksPerSm ? UB_PF_G2I_MINB : kBlocksPerSm) * waves;
  sh.blocks = (int)(blocksNeeded < persistent ? blocksNeeded : persistent);
  if (sh.blocks < 1) sh.blocks = 1;
  return sh;
}

}  // namespace

cudaError_t launchPfFrameletOrientations(const LsState &S, double *out, cudaStream_t stream) {
  int n = S.numFramelets + 2;
  pfFrameletOrientKernel<<<(n + 63) / 64, 64, 0, stream>>>(S, out);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchPfSensorPosition(const LsState &S, double time, double *xyz,
                                   cudaStream_t stream) {
  pfSensorPositionKernel<<<1, 1, 0, stream>>>(S, time, xyz);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchPfSensorState(const LsState &S, long n, const double *line,
                                const SensorStateOut &o, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  const long blocks = (n + 255) / 256;
  pfSensorStateKernel<<<(int)(blocks < 148 * 8 ? blocks : 148 * 8), 256, 0, stream>>>(S, n, line, o);
  countLaunch();
  return cudaGetLastError();
}

// USGSCSM_B200_PF_SPECIALISE=0: the generic kernels only
static bool pfSpecialise() {
  static const bool on = [] {
    const char *v = getenv("USGSCSM_B200_PF_SPECIALISE");
    return !(v && *v == '0');
  }();
  return on;
}

// see pfGroundToImagePoint
static bool pfFrameletNumbersExact(const LsState &S) {
  return S.frameletOrient && S.numFramelets >= 1 && S.summedFrameletHeight >= 1 &&
         (2.0 * S.numFramelets + 1.0) * S.summedFrameletHeight < 2147483648.0;
}

cudaError_t launchPfGroundToImage(const LsState &S, int smCount, long n, const double *x,
                                  const double *y, const double *z, double prec, double *line,
                                  double *samp, double *achieved, uint32_t *status,
                                  cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = pfShape(S, n, smCount, smemBytes, stage);
  WorkCounter wc;
  cudaError_t e = wc.acquire(stream, n);
  if (e != cudaSuccess) return e;
#define UB_PF_G2I(STAGE, DIST, ATT)                                                          \
  pfGroundToImageKernel<STAGE, DIST, ATT><<<sh.blocks, sh.threads, STAGE ? smemBytes : 0, stream>>>( \
      S, n, x, y, z, prec, line, samp, achieved, status, wc.ptr, grabFor(n))
#define UB_PF_G2I_ATT(STAGE, DIST)                \
  do {                                            \
    if (S.hasAttAdj) UB_PF_G2I(STAGE, DIST, 1);   \
    else UB_PF_G2I(STAGE, DIST, 0);               \
  } while (0)
#define UB_PF_G2I_DIST(STAGE)                                                             \
  do {                                                                                    \
    if (!pfSpecialise() || !pfFrameletNumbersExact(S)) UB_PF_G2I(STAGE, -1, -1);          \
    else if (S.fp.distortionType == DIST_RADIAL) UB_PF_G2I_ATT(STAGE, DIST_RADIAL);       \
    else if (S.fp.distortionType == DIST_TRANSVERSE && transverseCoeffsPlain(S.fp)) UB_PF_G2I_ATT(STAGE, DIST_TRANSVERSE); \
    else UB_PF_G2I(STAGE, -1, -1);                                                        \
  } while (0)
  if (stage)
    UB_PF_G2I_DIST(true);
  else
    UB_PF_G2I_DIST(false);
#undef UB_PF_G2I_DIST
#undef UB_PF_G2I_ATT
#undef UB_PF_G2I
  countLaunch();
  return wc.release();
}

cudaError_t launchPfImageToGround(const LsState &S, int smCount, long n, const double *line,
                                  const double *samp, const GridSpec *grid, const double *height,
                                  int heightStride, double prec, double *x, double *y, double *z,
                                  double *achieved, uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = pfShape(S, n, smCount, smemBytes, stage, true);
  GridSpec g = grid ? *grid : GridSpec{0, 0, 0, 0, 1, 0};
#define UB_PF_I2G(STAGE, GRID, DIST, ATT)                                                       \
  pfImageToGroundKernel<STAGE, GRID, DIST, ATT><<<sh.blocks, sh.threads, smemBytes, stream>>>(  \
      S, n, line, samp, g, height, heightStride, prec, x, y, z, achieved, status)
#define UB_PF_I2G_ATT(STAGE, GRID, DIST)                \
  do {                                                  \
    if (S.hasAttAdj) UB_PF_I2G(STAGE, GRID, DIST, 1);   \
    else UB_PF_I2G(STAGE, GRID, DIST, 0);               \
  } while (0)
#define UB_PF_I2G_DIST(STAGE, GRID)                                                               \
  do {                                                                                            \
    if (!pfSpecialise() || S.summedFrameletHeight < 1) UB_PF_I2G(STAGE, GRID, -1, -1);            \
    else if (S.fp.distortionType == DIST_RADIAL) UB_PF_I2G_ATT(STAGE, GRID, DIST_RADIAL);         \
    else if (S.fp.distortionType == DIST_TRANSVERSE && transverseCoeffsPlain(S.fp)) UB_PF_I2G_ATT(STAGE, GRID, DIST_TRANSVERSE); \
    else UB_PF_I2G(STAGE, GRID, -1, -1);                                                          \
  } while (0)
  if (stage && grid)
    UB_PF_I2G_DIST(true, true);
  else if (stage)
    UB_PF_I2G_DIST(true, false);
  else if (grid)
    UB_PF_I2G_DIST(false, true);
  else
    UB_PF_I2G_DIST(false, false);
#undef UB_PF_I2G_DIST
#undef UB_PF_I2G_ATT
#undef UB_PF_I2G
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchPfRemoteLocus(const LsState &S, int smCount, long n, const double *line,
                                const double *samp, const LocusOut &o, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = pfShape(S, n, smCount, smemBytes, stage);
  if (stage)
    pfRemoteLocusKernel<true><<<sh.blocks, sh.threads, smemBytes, stream>>>(S, n, line, samp, o);
  else
    pfRemoteLocusKernel<false><<<sh.blocks, sh.threads, 0, stream>>>(S, n, line, samp, o);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchGroundPartialsPf(const LsState &S, int smCount, long n, const double *x,
                                   const double *y, const double *z, double *partials,
                                   uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = pfShape(S, n, smCount, smemBytes, stage);
  WorkCounter wc;
  cudaError_t e = wc.acquire(stream, n);
  if (e != cudaSuccess) return e;
  if (stage)
    pfGroundPartialsKernel<true><<<sh.blocks, sh.threads, smemBytes, stream>>>(
        S, n, x, y, z, partials, status, wc.ptr, grabFor(n));
  else
    pfGroundPartialsKernel<false><<<sh.blocks, sh.threads, 0, stream>>>(
        S, n, x, y, z, partials, status, wc.ptr, grabFor(n));
  countLaunch();
  return wc.release();
}

cudaError_t launchPfProximateLocus(const LsState &S, int smCount, long n, const double *line,
                                   const double *samp, const double *x, const double *y,
                                   const double *z, double prec, const LocusOut &o,
                                   cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = pfShape(S, n, smCount, smemBytes, stage);
  if (stage)
    pfProximateLocusKernel<true><<<sh.blocks, sh.threads, smemBytes, stream>>>(S, n, line, samp, x, y,
                                                                               z, prec, o);
  else
    pfProximateLocusKernel<false><<<sh.blocks, sh.threads, 0, stream>>>(S, n, line, samp, x, y, z,
                                                                        prec, o);
  countLaunch();
  return cudaGetLastError();
}

}  // namespace ub200
