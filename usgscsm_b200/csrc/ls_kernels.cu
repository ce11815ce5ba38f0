// ls_kernels.cu -- line-scan groundToImage / imageToGround, one thread per point, FP64.
//
// Replaces the per-point calls UsgsAstroLsSensorModel::groundToImage
// (src/UsgsAstroLsSensorModel.cpp:757-890) and ::imageToGround (:980-1009).
//
// Layout: SoA double arrays in HBM, read/written once with streaming cache hints (a warp
// touches 32 consecutive doubles per array = two full 128-byte lines); the packed model
// state is a __grid_constant__ kernel parameter (constant bank); the ephemeris and
// attitude tables are staged once per block into shared memory when they fit, and the
// grid is persistent (a multiple of the SM count, grid-stride over points) so that the
// staging cost is paid once per resident block, not once per 256 points.
#include <cstdint>
#include <cstdlib>
#include <initializer_list>
#include <mutex>
#include <type_traits>

#include "geo_device.cuh"
#include "kernels.h"
#include "ls_common.cuh"
#include "work_queue.cuh"

namespace ub200 {

namespace {

constexpr int kThreads = 256;         // __launch_bounds__ cap
constexpr int kDefaultMinBlocks = 4;
constexpr int kDefaultThreads = 256;  // block size used unless overridden

// ---- groundToImage ---------------------------------------------------------------------
// One point of LS::groundToImage(groundPt, adj = 0, desiredPrecision, ...) LS:795-890.
//
// The reference evaluates calcDetectorLineErr (LS:3209-3231) at t = 0, t = 0.1 and at
// every secant iterate, then evaluates computeDetectorView once more at approxPt.line + t1.
// That last call sees exactly the time of the last residual evaluation (same sum
// approxPt.line + t1, same getImageTime), so its focal-plane point IS the one already
// computed; only applyDistortion is repeated, now with the caller's desiredPrecision
// instead of the default 1e-6 (hazard 3 in SURVEY.md §9).  The loop below therefore has a
// single inlined copy of the detector view and runs it 2 + k times.
template <bool ALWAYS_VEL, class TAB, class SPEC = GenericSpec, class DIV = ScaleDivision>
UB_DEV void lsGroundToImagePoint(const LsState &S, const TAB &T, double gx,
                                 double gy, double gz, double desiredPrecision, double &outLine,
                                 double &outSamp, double &achieved, uint32_t &status,
                                 const DIV &div = DIV()) {
  double approxLine, approxSamp;
  projectiveSeed<SPEC>(S, gx, gy, gz, approxLine, approxSamp);

  const FocalPlane &fp = S.fp;
  int count = 0;
  double t0 = 0.0, lineErr0 = 0.0;
  double t1 = 0.0, lineErr1;
  double focalX, focalY;
  double distortedFocalX, distortedFocalY;
  uint32_t distStatus;
  int stage = 0;  // 0: residual at t = 0; 1: residual at t1 (0.1, then secant iterates)
  uint32_t evals = 0;
  for (;;) {
    evals++;
    double timei = lsImageTime<SPEC>(S, approxLine + t1);
    detectorView<ALWAYS_VEL, TAB, SPEC, DIV>(S, T, timei, gx, gy, gz, focalX, focalY, div);
    distStatus = 0;
    applyDistortion<SPEC::kDist>(focalX, focalY, distortedFocalX, distortedFocalY, fp, 1.0E-6, distStatus);
    double detectorLine = fp.iTransL[0] + fp.iTransL[1] * distortedFocalX +
                          fp.iTransL[2] * distortedFocalY + fp.detectorLineOrigin -
                          fp.startingDetectorLine;
    detectorLine = divideBy(detectorLine, fp.lineSumming);
    if (stage == 0) {
      lineErr0 = detectorLine;
      t1 = 0.1;
      stage = 1;
      continue;
    }
    lineErr1 = detectorLine;
    // while (std::abs(lineErr1) > desiredPrecision && ++count < 15) { if (==) break; ... }
    if (!(fabs(lineErr1) > desiredPrecision)) break;
    if (!(++count < 15)) {
      status |= USGSCSM_B200_ST_ITERATION_CAP;
      break;
    }
    if (lineErr1 == lineErr0) {
      status |= USGSCSM_B200_ST_SEARCH_STALL;
      break;
    }
    double t2 = t1 - lineErr1 * (t1 - t0) / (lineErr1 - lineErr0);
    t0 = t1;
    lineErr0 = lineErr1;
    t1 = t2;
  }

  status |= evals << USGSCSM_B200_ST_EVALS_SHIFT;
  approxLine += t1;
  // Final projection: same focal-plane point, distortion at the caller's precision.  Only the
  // RADIAL and CAHVOR models read that precision (D:490, D:754; the others iterate to hard-coded
  // tolerances), so for every other model -- and for those two at the default 1e-6 -- this call
  // would recompute the last residual's distorted point bit for bit: it is reused.
  const int distType = SPEC::kDist >= 0 ? SPEC::kDist : fp.distortionType;
  if ((distType == DIST_RADIAL || distType == DIST_CAHVOR) && desiredPrecision != 1.0E-6) {
    distStatus = 0;
    applyDistortion<SPEC::kDist>(focalX, focalY, distortedFocalX, distortedFocalY, fp,
                                 desiredPrecision, distStatus);
  }
  status |= distStatus;
  double detectorLine =
      fp.iTransL[0] + fp.iTransL[1] * distortedFocalX + fp.iTransL[2] * distortedFocalY;
  double detectorSample =
      fp.iTransS[0] + fp.iTransS[1] * distortedFocalX + fp.iTransS[2] * distortedFocalY;
  double finalUpdate =
      divideBy(detectorLine + fp.detectorLineOrigin - fp.startingDetectorLine, fp.lineSumming);
  outLine = approxLine + finalUpdate;
  outSamp = divideBy(detectorSample + fp.detectorSampleOrigin - fp.startingDetectorSample,
                     fp.sampleSumming);
  achieved = finalUpdate;
  if ((desiredPrecision > 0.0) && (fabs(finalUpdate) > desiredPrecision))
    status |= USGSCSM_B200_ST_PRECISION_NOT_MET;
}

// Where the ground points of a launch come from: three SoA arrays, or the samples of a DEM
// window (the ground point of a sample is generated in the kernel: 2-8 bytes per point read
// instead of 24; usgscsm_b200_ground_to_image_dem_grid).
struct ArrayGround {
  const double *__restrict__ x, *__restrict__ y, *__restrict__ z;
  UB_DEV bool load(long i, double &gx, double &gy, double &gz) const {
    gx = __ldcs(x + i);
    gy = __ldcs(y + i);
    gz = __ldcs(z + i);
    return true;
  }
};
struct DemGround {
  DemWindow w;
  UB_DEV bool load(long i, double &gx, double &gy, double &gz) const {
    return demGroundPoint(w, i, gx, gy, gz);
  }
};

// Where the results of a launch go: the four output arrays of groundToImage, or -- for the
// per-parameter searches of computeAllSensorPartials (csm/RasterGM.cpp:101-140) -- straight into
// slot k of the point's list of partial derivatives, with the achieved precision and the
// warnings accumulated over the parameters (what a separate "step" kernel did before, at 0.3 ms
// of uncoalesced traffic per parameter).
struct PlainOut {
  double *__restrict__ line, *__restrict__ samp, *__restrict__ achieved;
  uint32_t *__restrict__ status;
  UB_DEV void store(long i, double l, double s, double a, uint32_t st) const {
    __stcs(line + i, l);
    __stcs(samp + i, s);
    if (achieved) __stcs(achieved + i, a);
    if (status) __stcs(status + i, st);
  }
};
struct PartialsOut {
  const double *__restrict__ l0, *__restrict__ s0;  // the image point the differences are taken from
  const uint32_t *__restrict__ st0;                 // its status (null when the caller supplied the point)
  double delta;
  double *__restrict__ partials;  // numParams x (dline, dsamp) per point
  int numParams, k;
  double *__restrict__ achieved;  // nullable: running maximum over the parameters
  uint32_t *__restrict__ status;  // running accumulation of the warning / error bits
  UB_DEV void store(long i, double l1, double s1, double a, uint32_t stP) const {
    const uint32_t kErr = USGSCSM_B200_ST_ERROR_ALGORITHM | USGSCSM_B200_ST_ERROR_UNKNOWN |
                          USGSCSM_B200_ST_ERROR_OTHER;
    const uint32_t kKeep = kErr | USGSCSM_B200_ST_NO_INTERSECTION | USGSCSM_B200_ST_PRECISION_NOT_MET;
    uint32_t acc = k == 0 ? (st0 ? st0[i] & kKeep : 0u) : status[i];
    double am = (k == 0 || !achieved) ? 0.0 : achieved[i];
    acc |= stP & kKeep;
    if (a > am) am = a;
    const double pl = (l1 - l0[i]) / delta;
    const double ps = (s1 - s0[i]) / delta;
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
};

template <class SPEC, class TAB, class SRC, class OUT>
UB_DEV void lsGroundToImageLoop(const LsState &S, const TAB &T, long n, const SRC &src,
                                double prec, const OUT &out,
                                unsigned long long *__restrict__ next, int grab) {
  // (reciprocals of m_halfTime / m_flyingHeight / m_halfSwath hoisted out of the loop --
  // HoistedScaleDivision -- measured 9 % SLOWER here: three more live doubles at 64 registers)
  using DIV = ScaleDivision;
  const DIV div(S);
  forEachPointDynamic(n, next, grab, [&](long i) {
    double gx, gy, gz;
    double l, s, a;
    uint32_t st = 0;
    if (src.load(i, gx, gy, gz)) {
      lsGroundToImagePoint<false, TAB, SPEC, DIV>(S, T, gx, gy, gz, prec, l, s, a, st, div);
      st |= imageRangeStatus(l, s, S.nLines, S.nSamples);
    } else {  // a nodata sample: nothing to project
      l = s = a = nan("");
      st = USGSCSM_B200_ST_DEM_NODATA | USGSCSM_B200_ST_NONFINITE;
    }
    out.store(i, l, s, a, st);
  });
}

// MINB = resident blocks of 256 threads per SM the register allocation is held to
// (2: 96+ registers, 3: 80, 4: 64).
//
// DIST >= 0 is the specialised kernel for the common case -- no parameter adjustments, one
// line-rate segment, distortion model DIST: the launcher has checked those facts of S, and
// making them compile-time facts (CommonSpec) removes every other branch (the eight other distortion
// models, the adjusted-position frame, the Euler-angle correction, the line-rate search) from
// the instruction stream.  Only ~1/6 of the generic kernel's instructions are ever executed
// for such a model, scattered over 85 KB of code, and ncu shows the search kernel waiting on
// instruction fetch (icc hit rate 88 %, "no instruction" 14 % of stall samples).
// DIST = -1 is the generic kernel.  Same arithmetic either way.
template <bool STAGE, int MINB, int DIST, int ADJ, class SRC, class OUT>
__global__ void __launch_bounds__(kThreads, MINB)
    lsGroundToImageKernel(const __grid_constant__ LsState S, long n,
                          const __grid_constant__ SRC src, double prec,
                          const __grid_constant__ OUT out,
                          unsigned long long *__restrict__ next, int grab) {
  extern __shared__ __align__(16) double smem[];
  using SPEC = typename std::conditional<(DIST >= 0), CommonSpec<DIST, ADJ>, SearchSpec>::type;
  if (STAGE) {
    StagedTables T = stageOrientationTables(S, specHasAdj<SPEC>(S), smem);
    lsGroundToImageLoop<SPEC>(S, T, n, src, prec, out, next, grab);
  } else {
    OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
    lsGroundToImageLoop<SPEC>(S, T, n, src, prec, out, next, grab);
  }
}

// ---- imageToGround -----------------------------------------------------------------------
// One point of LS::imageToGround LS:980-1009 = losToEcf LS:2209-2305 + losEllipsoidIntersect.
template <class TAB>
UB_DEV void lsImageToGroundPoint(const LsState &S, const TAB &T, double line,
                                 double sample, double height, double desiredPrecision,
                                 double &x, double &y, double &z, double &achieved,
                                 uint32_t &status) {
  double time = lsImageTime(S, line);
  double pos[3], vel[3];
  adjSensorPosVel<false>(S, T, time, pos, vel);
  // The detector is one row: the focal-plane affine is fed line = 0.0 (LS:2242-2246).
  double distortedFocalPlaneX, distortedFocalPlaneY;
  distortedFocalPlaneCoordinates(0.0, sample, S.fp, distortedFocalPlaneX, distortedFocalPlaneY);
  double undistortedFocalPlaneX, undistortedFocalPlaneY;
  removeDistortion(distortedFocalPlaneX, distortedFocalPlaneY, undistortedFocalPlaneX,
                   undistortedFocalPlaneY, S.fp, 1.0E-6, status);
  double cameraLook[3];
  cameraLookVector(undistortedFocalPlaneX, undistortedFocalPlaneY, S.fp.zDirection,
                   S.fp.focalLength * (1 - param(S, 15) / S.halfSwath), cameraLook);
  double bodyLook[3];
  cameraLookToBody(S, T, time, cameraLook, bodyLook);
  lsEllipsoidIntersect(S.ell, height, pos, bodyLook, desiredPrecision, x, y, z, achieved, status);
  if ((desiredPrecision > 0.0) && (achieved > desiredPrecision))
    status |= USGSCSM_B200_ST_PRECISION_NOT_MET;
  if (status & USGSCSM_B200_ST_ERROR_ALGORITHM) {
    // the scalar call throws before producing anything
    status = USGSCSM_B200_ST_ERROR_ALGORITHM;
    x = y = z = achieved = nan("");
  }
  if (!finite3(x, y, z)) status |= USGSCSM_B200_ST_NONFINITE;
}

// ---- dense-grid imageToGround with row / column tables -------------------------------------
// On a dense grid everything the reference recomputes per pixel from the LINE alone (image
// time, Lagrange position, Lagrange quaternion, rotation matrix, attitude correction) is
// shared by a whole image row, and everything it recomputes from the SAMPLE alone (focal
// plane affine, distortion removal, unit look vector) by a whole column.  A set-up kernel
// evaluates both once per row / column with the per-point device functions (same arithmetic
// => same bits); the per-pixel kernel is left with the two rotations, the ellipsoid
// intersection and the elevation iteration.
constexpr int kRowDoubles = 21;  // pos[3], camToBody[9], attCorr[9]
constexpr int kColDoubles = 4;   // cameraLook[3], status bits of removeDistortion

__global__ void lsGridTablesKernel(const __grid_constant__ LsState S, GridSpec g, long nRows,
                                   long nCols, double *__restrict__ rowTab,
                                   double *__restrict__ colTab) {
  long k = (long)blockIdx.x * blockDim.x + threadIdx.x;
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  if (k < nRows) {
    double line = g.line0 + (double)(g.row0 + k) * g.dline;
    double time = lsImageTime(S, line);
    double pos[3], vel[3], m[9], a[9];
    adjSensorPosVel<false>(S, T, time, pos, vel);
    cameraToBodyMatrix(S, T, time, m);
    for (int i = 0; i < 9; i++) a[i] = 0.0;
    if (S.hasAttAdj) attitudeCorrection(S, time, a);
    double *e = rowTab + k * kRowDoubles;
    for (int i = 0; i < 3; i++) e[i] = pos[i];
    for (int i = 0; i < 9; i++) e[3 + i] = m[i];
    for (int i = 0; i < 9; i++) e[12 + i] = a[i];
  } else if (k < nRows + nCols) {
    long c = k - nRows;
    double sample = g.samp0 + (double)c * g.dsamp;
    uint32_t st = 0;
    double dx, dy, ux, uy, look[3];
    distortedFocalPlaneCoordinates(0.0, sample, S.fp, dx, dy);
    removeDistortion(dx, dy, ux, uy, S.fp, 1.0E-6, st);
    cameraLookVector(ux, uy, S.fp.zDirection, S.fp.focalLength * (1 - param(S, 15) / S.halfSwath),
                     look);
    double *e = colTab + c * kColDoubles;
    e[0] = look[0];
    e[1] = look[1];
    e[2] = look[2];
    e[3] = (double)st;
  }
}

// (Two pixels per thread with 128-bit accesses -- what pays in the frame kernels -- measured
// 74.3 against 73.9 Gpt/s here at its best register budget: not kept.)
template <int MINB>
__global__ void __launch_bounds__(kThreads, MINB)
    lsImageToGroundGridKernel(const __grid_constant__ LsState S, long n, long nSamps,
                              const double *__restrict__ rowTab, const double *__restrict__ colTab,
                              const double *__restrict__ height, int heightStride, double prec,
                              double *__restrict__ x, double *__restrict__ y, double *__restrict__ z,
                              double *__restrict__ achieved, uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  // the height is the one input that comes from DRAM: the next pass's is requested before this
  // pass's arithmetic starts (ncu: long-scoreboard stalls 3.1 per issue without it)
  double hNext = (i < n) ? (heightStride ? __ldcs(height + i) : __ldg(height)) : 0.0;
  // (row, column) of the thread's pixel, advanced by the stride's (rows, columns) instead of a
  // 64-bit division per pixel
  long r = i / nSamps, c = i - r * nSamps;
  const long strideRows = stride / nSamps, strideCols = stride - strideRows * nSamps;
  for (; i < n; i += stride, r += strideRows, c += strideCols) {
    if (c >= nSamps) {
      c -= nSamps;
      r++;
    }
    const double *re = rowTab + r * kRowDoubles;
    const double2 *ce = reinterpret_cast<const double2 *>(colTab + c * kColDoubles);  // 16-byte aligned
    const double h = hNext;
    if (heightStride && i + stride < n) hNext = __ldcs(height + i + stride);
    const double2 c01 = __ldg(ce), c23 = __ldg(ce + 1);
    double cameraLook[3] = {c01.x, c01.y, c23.x};
    uint32_t st = (uint32_t)c23.y;
    double pos[3] = {__ldg(re), __ldg(re + 1), __ldg(re + 2)};
    double m[9];
#pragma unroll
    for (int k = 0; k < 9; k++) m[k] = __ldg(re + 3 + k);
    double corrected[3];
    if (S.hasAttAdj) {
      double a[9];
#pragma unroll
      for (int k = 0; k < 9; k++) a[k] = __ldg(re + 12 + k);
      matVec(a, cameraLook, corrected);
    } else {
      corrected[0] = cameraLook[0];
      corrected[1] = cameraLook[1];
      corrected[2] = cameraLook[2];
    }
    double bodyLook[3];
    matVec(m, corrected, bodyLook);
    double px, py, pz, ach;
    lsEllipsoidIntersect(S.ell, h, pos, bodyLook, prec, px, py, pz, ach, st);
    if ((prec > 0.0) && (ach > prec)) st |= USGSCSM_B200_ST_PRECISION_NOT_MET;
    if (st & USGSCSM_B200_ST_ERROR_ALGORITHM) {
      st = USGSCSM_B200_ST_ERROR_ALGORITHM;
      px = py = pz = ach = nan("");
    }
    if (!finite3(px, py, pz)) st |= USGSCSM_B200_ST_NONFINITE;
    __stcs(x + i, px);
    __stcs(y + i, py);
    __stcs(z + i, pz);
    if (achieved) __stcs(achieved + i, ach);
    if (status) __stcs(status + i, st);
  }
}

template <bool GRID, class TAB>
UB_DEV void lsImageToGroundLoop(const LsState &S, const TAB &T, long n, const double *__restrict__ line,
                                const double *__restrict__ samp, const GridSpec &g,
                                const double *__restrict__ height, int heightStride, double prec,
                                double *__restrict__ x, double *__restrict__ y, double *__restrict__ z,
                                double *__restrict__ achieved, uint32_t *__restrict__ status) {
  // static grid-stride split: no search in this kernel, every point costs the same, and the
  // dynamic queue measured 2 % slower here
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double l, s;
    if (GRID) {
      long r = i / g.nSamps;
      long c = i - r * g.nSamps;
      l = g.line0 + (double)(g.row0 + r) * g.dline;
      s = g.samp0 + (double)c * g.dsamp;
    } else {
      l = __ldcs(line + i);
      s = __ldcs(samp + i);
    }
    double h = heightStride ? __ldcs(height + i) : __ldg(height);
    double px, py, pz, a;
    uint32_t st = 0;
    lsImageToGroundPoint(S, T, l, s, h, prec, px, py, pz, a, st);
    __stcs(x + i, px);
    __stcs(y + i, py);
    __stcs(z + i, pz);
    if (achieved) __stcs(achieved + i, a);
    if (status) __stcs(status + i, st);
  }
}

// GRID = true: pixel coordinates generated from the point index (dense grid);
// GRID = false: explicit (line, samp) arrays.
template <bool STAGE, bool GRID, int MINB>
__global__ void __launch_bounds__(kThreads, MINB)
    lsImageToGroundKernel(const __grid_constant__ LsState S, long n, const double *__restrict__ line,
                          const double *__restrict__ samp, GridSpec g,
                          const double *__restrict__ height, int heightStride, double prec,
                          double *__restrict__ x, double *__restrict__ y, double *__restrict__ z,
                          double *__restrict__ achieved, uint32_t *__restrict__ status) {
  extern __shared__ __align__(16) double smem[];
  if (STAGE) {
    StagedTables T = stageOrientationTables(S, S.hasAdj != 0, smem);
    lsImageToGroundLoop<GRID>(S, T, n, line, samp, g, height, heightStride, prec, x, y, z, achieved,
                              status);
  } else {
    OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
    lsImageToGroundLoop<GRID>(S, T, n, line, samp, g, height, heightStride, prec, x, y, z, achieved,
                              status);
  }
}

// ---- computeGroundPartials ----------------------------------------------------------------
// LS:1299-1325: groundToImage (default precision 0.001) at the point and at the point moved by
// m_gsd along X, Y, Z; forward differences.  One kernel, four searches per point.
// DIST >= 0: the specialised search (one line-rate segment, projective seed, distortion DIST;
// parameter adjustments read from the state), as in lsGroundToImageKernel.
template <bool STAGE, int DIST>
__global__ void __launch_bounds__(kThreads, kDefaultMinBlocks)
    lsGroundPartialsKernel(const __grid_constant__ LsState S, long n, const double *__restrict__ x,
                           const double *__restrict__ y, const double *__restrict__ z,
                           double *__restrict__ partials, uint32_t *__restrict__ status,
                           unsigned long long *__restrict__ next, int grab) {
  extern __shared__ __align__(16) double smem[];
  using SPEC = typename std::conditional<(DIST >= 0), CommonSpec<DIST, -1>, SearchSpec>::type;
  StagedTables TS{};
  OrientationTables TG{S.positions.values, S.velocities.values, S.quaternions.values};
  if (STAGE) TS = stageOrientationTables(S, S.hasAdj != 0, smem);
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
      if (STAGE)
        lsGroundToImagePoint<false, StagedTables, SPEC>(S, TS, q[0], q[1], q[2], 0.001, l[k], s[k], a, sk);
      else
        lsGroundToImagePoint<false, OrientationTables, SPEC>(S, TG, q[0], q[1], q[2], 0.001, l[k], s[k], a, sk);
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

// ---- imageToRemoteImagingLocus -----------------------------------------------------------
// LS:1254-1282: losToEcf (LS:2209-2305) at the pixel; point = sensor position, direction =
// MINUS the body-fixed look vector losToEcf returns (the reference negates it, LS:1267-1270);
// achievedPrecision = 0.  Callers ray-trace DEMs from this.
template <class TAB>
UB_DEV void lsRemoteLocusPoint(const LsState &S, const TAB &T, double line, double sample,
                               double pos[3], double dir[3], uint32_t &status) {
  double time = lsImageTime(S, line);
  double vel[3];
  adjSensorPosVel<false>(S, T, time, pos, vel);
  double distortedFocalPlaneX, distortedFocalPlaneY;
  distortedFocalPlaneCoordinates(0.0, sample, S.fp, distortedFocalPlaneX, distortedFocalPlaneY);
  double undistortedFocalPlaneX, undistortedFocalPlaneY;
  removeDistortion(distortedFocalPlaneX, distortedFocalPlaneY, undistortedFocalPlaneX,
                   undistortedFocalPlaneY, S.fp, 1.0E-6, status);
  double cameraLook[3];
  cameraLookVector(undistortedFocalPlaneX, undistortedFocalPlaneY, S.fp.zDirection,
                   S.fp.focalLength * (1 - param(S, 15) / S.halfSwath), cameraLook);
  double bodyLook[3];
  cameraLookToBody(S, T, time, cameraLook, bodyLook);
  dir[0] = -bodyLook[0];
  dir[1] = -bodyLook[1];
  dir[2] = -bodyLook[2];
  if (status & USGSCSM_B200_ST_ERROR_ALGORITHM) {
    status = USGSCSM_B200_ST_ERROR_ALGORITHM;
    pos[0] = pos[1] = pos[2] = dir[0] = dir[1] = dir[2] = nan("");
  }
  if (!(finite3(pos[0], pos[1], pos[2]) && finite3(dir[0], dir[1], dir[2])))
    status |= USGSCSM_B200_ST_NONFINITE;
}

template <bool STAGE>
__global__ void __launch_bounds__(kThreads)
    lsRemoteLocusKernel(const __grid_constant__ LsState S, long n, const double *__restrict__ line,
                        const double *__restrict__ samp, LocusOut o) {
  extern __shared__ __align__(16) double smem[];
  StagedTables TS{};
  OrientationTables TG{S.positions.values, S.velocities.values, S.quaternions.values};
  if (STAGE) TS = stageOrientationTables(S, S.hasAdj != 0, smem);
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double pos[3], dir[3];
    uint32_t st = 0;
    if (STAGE)
      lsRemoteLocusPoint(S, TS, __ldcs(line + i), __ldcs(samp + i), pos, dir, st);
    else
      lsRemoteLocusPoint(S, TG, __ldcs(line + i), __ldcs(samp + i), pos, dir, st);
    storeLocus(o, i, pos, dir, 0.0, st);
  }
}

// ---- imageToProximateImagingLocus ----------------------------------------------------------
template <bool STAGE>
__global__ void __launch_bounds__(kThreads, 3)
    lsProximateLocusKernel(const __grid_constant__ LsState S, long n, const double *__restrict__ line,
                           const double *__restrict__ samp, const double *__restrict__ x,
                           const double *__restrict__ y, const double *__restrict__ z, double prec,
                           LocusOut o) {
  extern __shared__ __align__(16) double smem[];
  StagedTables TS{};
  OrientationTables TG{S.positions.values, S.velocities.values, S.quaternions.values};
  if (STAGE) TS = stageOrientationTables(S, S.hasAdj != 0, smem);
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double l = __ldcs(line + i), sm = __ldcs(samp + i);
    double pos[3], dir[3], a;
    uint32_t st = 0;
    proximateLocusFromImageToGround(
        S.ell, S.gsd, __ldcs(x + i), __ldcs(y + i), __ldcs(z + i), prec,
        [&](double h, double g[3], double &ach, uint32_t &s) {
          if (STAGE)
            lsImageToGroundPoint(S, TS, l, sm, h, prec, g[0], g[1], g[2], ach, s);
          else
            lsImageToGroundPoint(S, TG, l, sm, h, prec, g[0], g[1], g[2], ach, s);
        },
        pos, dir, a, st);
    storeLocus(o, i, pos, dir, a, st);
  }
}

// ---- model set-up helpers -------------------------------------------------------------------
// The reference derives a few per-model inputs of the hot path at load time with its own
// per-point routines (updateState LS:649-735, createProjectiveApproximation LS:2745-2803).
// The engine has no CPU copy of that math, so the loader runs these two small kernels.

// getSensorPosition(ImageCoord) LS:1580-1611: getAdjSensorPosVel with calc_vel = true.
__global__ void lsSensorPositionKernel(const __grid_constant__ LsState S, int n,
                                       const double *__restrict__ line, double *__restrict__ xyz) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  double pos[3], vel[3];
  adjSensorPosVel<true>(S, T, lsImageTime(S, line[i]), pos, vel);
  xyz[3 * i + 0] = pos[0];
  xyz[3 * i + 1] = pos[1];
  xyz[3 * i + 2] = pos[2];
}

// getImageTime LS:1543-1568, getSensorPosition(ImageCoord) LS:1580-1611 (calc_vel = false: the
// nominal position when no parameter is set), getSensorVelocity(ImageCoord) LS:1623-1653
// (calc_vel = true), per image point.
__global__ void lsSensorStateKernel(const __grid_constant__ LsState S, long n,
                                    const double *__restrict__ line, SensorStateOut o) {
  OrientationTables T{S.positions.values, S.velocities.values, S.quaternions.values};
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double time = o.atTime ? line[i] : lsImageTime(S, line[i]);
    double pos[3], vel[3], posV[3], unused[3];
    adjSensorPosVel<false>(S, T, time, pos, unused);
    adjSensorPosVel<true>(S, T, time, posV, vel);
    storeSensorState(o, i, time, pos, vel, 0u);
  }
}

// computeEllipsoidElevation U:819-868 for n points (xyz interleaved).
__global__ void ellipsoidElevationKernel(Ellipsoid ell, int n, const double *__restrict__ xyz,
                                         double prec, double *__restrict__ h) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  h[i] = ellipsoidElevation(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], ell, prec);
}

}  // namespace

// ---- launchers ---------------------------------------------------------------------------
static int envInt(const char *name, int def) {
  const char *v = getenv(name);
  if (!v || !*v) return def;
  int x = atoi(v);
  return x > 0 ? x : def;
}

// Launch shape: persistent grid of smCount x blocksPerSm blocks.  The two tuning knobs can
// be overridden from the environment for experiments (USGSCSM_B200_LS_THREADS,
// USGSCSM_B200_LS_BLOCKS_PER_SM); the defaults are the measured best on B200.
static LaunchShape shapeFor(const LsState &S, long n, int smCount, size_t &smemBytes, bool &stage,
                            int blocksPerSm) {
  static const int threads = envInt("USGSCSM_B200_LS_THREADS", kDefaultThreads);
  static const int perSm = envInt("USGSCSM_B200_LS_BLOCKS_PER_SM", 0);
  size_t bytes = orientationTableBytes(S, S.hasAdj != 0);
  stage = bytes <= kMaxStagedTableBytes;
  smemBytes = stage ? bytes : 0;
  LaunchShape sh;
  sh.threads = threads > kThreads ? kThreads : threads;
  if (perSm > 0) blocksPerSm = perSm;
  else blocksPerSm = blocksPerSm * kThreads / sh.threads;
  long blocksNeeded = (n + sh.threads - 1) / sh.threads;
  long persistent = (long)smCount * blocksPerSm;
  sh.blocks = (int)(blocksNeeded < persistent ? blocksNeeded : persistent);
  if (sh.blocks < 1) sh.blocks = 1;
  return sh;
}

template <class SRC, class OUT>
static cudaError_t launchLsGroundToImageFrom(const LsState &S, int smCount, long n, const SRC &src,
                                             double prec, const OUT &out, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = shapeFor(S, n, smCount, smemBytes, stage, kDefaultMinBlocks);
  // work counter of this launch: stream-ordered scratch, so that launches in flight on
  // different streams never share one
  const int grab = grabFor(n);
  WorkCounter wc;
  cudaError_t e = wc.acquire(stream, n);
  if (e != cudaSuccess) return e;
  unsigned long long *next = wc.ptr;
#define UB_LAUNCH_G2I(STAGE, DIST, ADJ)                                                       \
  do {                                                                                        \
    if (smemBytes > 48 * 1024)                                                                \
      cudaFuncSetAttribute(lsGroundToImageKernel<STAGE, kDefaultMinBlocks, DIST, ADJ, SRC, OUT>, \
                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);      \
    lsGroundToImageKernel<STAGE, kDefaultMinBlocks, DIST, ADJ, SRC, OUT>                      \
        <<<sh.blocks, sh.threads, smemBytes, stream>>>(S, n, src, prec, out, next, grab);     \
  } while (0)
#define UB_LAUNCH_G2I_ADJ(STAGE, DIST)                                                        \
  do {                                                                                        \
    if (adj == 0) UB_LAUNCH_G2I(STAGE, DIST, 0);                                              \
    else if (adj == 1) UB_LAUNCH_G2I(STAGE, DIST, 1);                                         \
    else if (adj == 2) UB_LAUNCH_G2I(STAGE, DIST, 2);                                         \
    else UB_LAUNCH_G2I(STAGE, DIST, 3);                                                       \
  } while (0)
#define UB_LAUNCH_G2I_DIST(STAGE)                                                             \
  do {                                                                                        \
    if (!common) UB_LAUNCH_G2I(STAGE, -1, 0);                                                 \
    else if (S.fp.distortionType == 4) UB_LAUNCH_G2I_ADJ(STAGE, 4); /* LROLROCNAC */          \
    else if (S.fp.distortionType == 0) UB_LAUNCH_G2I_ADJ(STAGE, 0); /* RADIAL */              \
    else if (S.fp.distortionType == 1 && transverseCoeffsPlain(S.fp)) UB_LAUNCH_G2I_ADJ(STAGE, 1); /* TRANSVERSE */          \
    else UB_LAUNCH_G2I(STAGE, -1, 0);                                                         \
  } while (0)
  // USGSCSM_B200_LS_SPECIALISE: 1 = specialise (default), 2 = only models without parameter
  // adjustments, 3 = never
  static const int specialise = envInt("USGSCSM_B200_LS_SPECIALISE", 1);
  const int adj = (S.hasAdj ? 1 : 0) | (S.hasAttAdj ? 2 : 0);
  const bool common = (specialise == 1 || (specialise == 2 && adj == 0)) && S.numIntTimes == 1 &&
                      S.useApproxInitTrans;
  if (stage)
    UB_LAUNCH_G2I_DIST(true);
  else
    UB_LAUNCH_G2I_DIST(false);
#undef UB_LAUNCH_G2I_ADJ
#undef UB_LAUNCH_G2I_DIST
#undef UB_LAUNCH_G2I
  countLaunch();
  return wc.release();
}

cudaError_t launchLsGroundToImage(const LsState &S, int smCount, long n, const double *x,
                                  const double *y, const double *z, double prec, double *line,
                                  double *samp, double *achieved, uint32_t *status,
                                  cudaStream_t stream) {
  return launchLsGroundToImageFrom(S, smCount, n, ArrayGround{x, y, z}, prec,
                                   PlainOut{line, samp, achieved, status}, stream);
}

// One parameter of computeAllSensorPartials: the search with the adjusted state S, its result
// differenced against (l0, s0) into slot k of the point's list.
cudaError_t launchLsSensorPartialStep(const LsState &S, int smCount, long n, const double *x,
                                      const double *y, const double *z, double prec, const double *l0,
                                      const double *s0, const uint32_t *st0, double delta,
                                      double *partials, int numParams, int k, double *achieved,
                                      uint32_t *status, cudaStream_t stream) {
  return launchLsGroundToImageFrom(S, smCount, n, ArrayGround{x, y, z}, prec,
                                   PartialsOut{l0, s0, st0, delta, partials, numParams, k, achieved, status},
                                   stream);
}

// groundToImage of the ground points of a DEM window, generated in the kernel.
cudaError_t launchLsGroundToImageDem(const LsState &S, int smCount, long n, const DemWindow &w,
                                     double prec, double *line, double *samp, double *achieved,
                                     uint32_t *status, cudaStream_t stream) {
  return launchLsGroundToImageFrom(S, smCount, n, DemGround{w}, prec,
                                   PlainOut{line, samp, achieved, status}, stream);
}

cudaError_t launchLsImageToGround(const LsState &S, int smCount, long n, const double *line,
                                  const double *samp, const GridSpec *grid, const double *height,
                                  int heightStride, double prec, double *x, double *y, double *z,
                                  double *achieved, uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  static const int useTables = envInt("USGSCSM_B200_LS_GRID_TABLES", 1);
  if (grid && grid->nSamps > 0 && n % grid->nSamps == 0 && useTables == 1) {
    // row / column tables in stream-ordered scratch, then the per-pixel kernel
    const long nCols = grid->nSamps, nRows = n / nCols;
    double *scratch = nullptr;
    cudaError_t e = cudaMallocAsync(&scratch, (nRows * kRowDoubles + nCols * kColDoubles) * sizeof(double), stream);
    if (e != cudaSuccess) return e;
    // column entries first: they are read with 128-bit loads and need the 16-byte boundary
    double *colTab = scratch, *rowTab = scratch + nCols * kColDoubles;
    long items = nRows + nCols;
    lsGridTablesKernel<<<(unsigned)((items + 127) / 128), 128, 0, stream>>>(S, *grid, nRows, nCols, rowTab, colTab);
    countLaunch();
    // register budget of the per-pixel kernel: resident blocks of 256 threads per SM
    // (USGSCSM_B200_LS_GRID_MINB; the default is the measured best on B200)
    static const int gridMinb = envInt("USGSCSM_B200_LS_GRID_MINB", 4);
    const int mb = gridMinb < 2 ? 2 : (gridMinb > 6 ? 6 : gridMinb);
    long blocksNeeded = (n + kThreads - 1) / kThreads;
    const long persistent = (long)smCount * mb;
    int blocks = (int)(blocksNeeded < persistent ? blocksNeeded : persistent);
#define UB_LAUNCH_GRID(MB)                                                                       \
  lsImageToGroundGridKernel<MB><<<blocks, kThreads, 0, stream>>>(S, n, nCols, rowTab, colTab, height, \
                                                                 heightStride, prec, x, y, z, achieved, status)
    if (mb == 2) UB_LAUNCH_GRID(2);
    else if (mb == 3) UB_LAUNCH_GRID(3);
    else if (mb == 4) UB_LAUNCH_GRID(4);
    else if (mb == 5) UB_LAUNCH_GRID(5);
    else UB_LAUNCH_GRID(6);
#undef UB_LAUNCH_GRID
    countLaunch();
    e = cudaGetLastError();
    cudaError_t e2 = cudaFreeAsync(scratch, stream);
    return e != cudaSuccess ? e : e2;
  }
  size_t smemBytes;
  bool stage;
  static const int minb = envInt("USGSCSM_B200_LS_I2G_MINB", 4);
  LaunchShape sh = shapeFor(S, n, smCount, smemBytes, stage, minb);
  GridSpec g = grid ? *grid : GridSpec{0, 0, 0, 0, 1, 0};
#define UB_LAUNCH_I2G_M(STAGE, GRID, MINB)                                                   \
  do {                                                                                       \
    if (smemBytes > 48 * 1024)                                                               \
      cudaFuncSetAttribute(lsImageToGroundKernel<STAGE, GRID, MINB>,                         \
                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);     \
    lsImageToGroundKernel<STAGE, GRID, MINB><<<sh.blocks, sh.threads, smemBytes, stream>>>(  \
        S, n, line, samp, g, height, heightStride, prec, x, y, z, achieved, status);         \
  } while (0)
#define UB_LAUNCH_I2G(STAGE, GRID)                                                           \
  do {                                                                                       \
    if (minb >= 4) UB_LAUNCH_I2G_M(STAGE, GRID, 4);                                          \
    else if (minb == 3) UB_LAUNCH_I2G_M(STAGE, GRID, 3);                                     \
    else UB_LAUNCH_I2G_M(STAGE, GRID, 2);                                                    \
  } while (0)
  if (stage && grid)
    UB_LAUNCH_I2G(true, true);
  else if (stage)
    UB_LAUNCH_I2G(true, false);
  else if (grid)
    UB_LAUNCH_I2G(false, true);
  else
    UB_LAUNCH_I2G(false, false);
#undef UB_LAUNCH_I2G
#undef UB_LAUNCH_I2G_M
  countLaunch();
  return cudaGetLastError();
}

// ---- dense-grid imageToGround without any allocation ----------------------------------------
// The row / column tables of launchLsImageToGround's dense-grid path live in stream-ordered
// scratch (cudaMallocAsync).  Callers that may not allocate (CUDA graph capture, serving loops)
// get the same two kernels over a file-scope array instead: every GPU that loads the library
// holds kStaticGridScratchDoubles * 8 bytes (16 MiB) until the process exits.  Every call shares
// the array and it keeps the last call's tables; lsGridTablesKernel writes every entry the
// per-pixel kernel reads (all rows, all columns of THIS call), so nothing has to start at zero.
// Two calls whose kernels overlap on one GPU (different streams, different host threads) race.
static __device__ __align__(256) double gStaticGridScratch[kStaticGridScratchDoubles];

// In host code the array's name is a stand-in, not a device address: the address is asked for
// once per device, with that device current.
static cudaError_t staticGridScratch(double **ptr) {
  constexpr int kMaxDevices = 64;
  static double *perDevice[kMaxDevices] = {};
  static std::mutex mu;
  int device = -1;
  cudaError_t e = cudaGetDevice(&device);
  if (e != cudaSuccess) return e;
  if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(mu);
  if (!perDevice[device]) {
    double *p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), gStaticGridScratch);
    if (e != cudaSuccess) return e;
    if (!p) return cudaErrorInvalidSymbol;
    perDevice[device] = p;
  }
  *ptr = perDevice[device];
  return cudaSuccess;
}

// The caller has checked nRows * kRowDoubles + nCols * kColDoubles <= kStaticGridScratchDoubles
// (staticGridScratchFits) and made the model's device current.
bool staticGridScratchFits(long nRows, long nCols) {
  return nRows >= 0 && nCols >= 0 && nRows <= kStaticGridScratchDoubles && nCols <= kStaticGridScratchDoubles &&
         nRows * kRowDoubles + nCols * kColDoubles <= (long)kStaticGridScratchDoubles;
}

cudaError_t launchLsImageToGroundGridStatic(const LsState &S, int smCount, const GridSpec &grid, long nRows,
                                            const double *height, int heightStride, double prec,
                                            double *x, double *y, double *z, double *achieved,
                                            uint32_t *status, cudaStream_t stream) {
  const long nCols = grid.nSamps, n = nRows * nCols;
  if (n <= 0) return cudaSuccess;
  if (!staticGridScratchFits(nRows, nCols)) return cudaErrorInvalidValue;
  double *scratch = nullptr;
  cudaError_t e = staticGridScratch(&scratch);
  if (e != cudaSuccess) return e;
  double *colTab = scratch, *rowTab = scratch + nCols * kColDoubles;  // 128-bit loads of column entries
  const long items = nRows + nCols;
  lsGridTablesKernel<<<(unsigned)((items + 127) / 128), 128, 0, stream>>>(S, grid, nRows, nCols, rowTab, colTab);
  countLaunch();
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const long blocksNeeded = (n + kThreads - 1) / kThreads;
  const long persistent = (long)smCount * 4;
  const int blocks = (int)(blocksNeeded < persistent ? blocksNeeded : persistent);
  lsImageToGroundGridKernel<4><<<blocks, kThreads, 0, stream>>>(S, n, nCols, rowTab, colTab, height, heightStride,
                                                               prec, x, y, z, achieved, status);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchLsSensorPosition(const LsState &S, int n, const double *line, double *xyz,
                                   cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  lsSensorPositionKernel<<<(n + 63) / 64, 64, 0, stream>>>(S, n, line, xyz);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchEllipsoidElevation(const Ellipsoid &ell, int n, const double *xyz, double prec,
                                     double *h, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  ellipsoidElevationKernel<<<(n + 63) / 64, 64, 0, stream>>>(ell, n, xyz, prec, h);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchLsRemoteLocus(const LsState &S, int smCount, long n, const double *line,
                                const double *samp, const LocusOut &o, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = shapeFor(S, n, smCount, smemBytes, stage, kBlocksPerSm);
  if (stage) {
    if (smemBytes > 48 * 1024)
      cudaFuncSetAttribute(lsRemoteLocusKernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)smemBytes);
    lsRemoteLocusKernel<true><<<sh.blocks, sh.threads, smemBytes, stream>>>(S, n, line, samp, o);
  } else {
    lsRemoteLocusKernel<false><<<sh.blocks, sh.threads, 0, stream>>>(S, n, line, samp, o);
  }
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchLsSensorState(const LsState &S, long n, const double *line,
                                const SensorStateOut &o, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  const long blocks = (n + 255) / 256;
  lsSensorStateKernel<<<(int)(blocks < 148 * 8 ? blocks : 148 * 8), 256, 0, stream>>>(S, n, line, o);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchGroundPartialsLs(const LsState &S, int smCount, long n, const double *x,
                                   const double *y, const double *z, double *partials,
                                   uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = shapeFor(S, n, smCount, smemBytes, stage, kDefaultMinBlocks);
  WorkCounter wc;
  cudaError_t e = wc.acquire(stream, n);
  if (e != cudaSuccess) return e;
#define UB_LAUNCH_GP(STAGE, DIST)                                                              \
  do {                                                                                        \
    if (smemBytes > 48 * 1024)                                                                \
      cudaFuncSetAttribute(lsGroundPartialsKernel<STAGE, DIST>,                               \
                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes);      \
    lsGroundPartialsKernel<STAGE, DIST><<<sh.blocks, sh.threads, smemBytes, stream>>>(        \
        S, n, x, y, z, partials, status, wc.ptr, grabFor(n));                                 \
  } while (0)
#define UB_LAUNCH_GP_DIST(STAGE)                                                              \
  do {                                                                                        \
    if (!common) UB_LAUNCH_GP(STAGE, -1);                                                     \
    else if (S.fp.distortionType == 4) UB_LAUNCH_GP(STAGE, 4);                                \
    else if (S.fp.distortionType == 0) UB_LAUNCH_GP(STAGE, 0);                                \
    else if (S.fp.distortionType == 1 && transverseCoeffsPlain(S.fp)) UB_LAUNCH_GP(STAGE, 1);                                \
    else UB_LAUNCH_GP(STAGE, -1);                                                             \
  } while (0)
  static const int specialise = envInt("USGSCSM_B200_LS_SPECIALISE", 1);
  const bool common = specialise == 1 && S.numIntTimes == 1 && S.useApproxInitTrans;
  if (stage)
    UB_LAUNCH_GP_DIST(true);
  else
    UB_LAUNCH_GP_DIST(false);
#undef UB_LAUNCH_GP_DIST
#undef UB_LAUNCH_GP
  countLaunch();
  return wc.release();
}

cudaError_t launchLsProximateLocus(const LsState &S, int smCount, long n, const double *line,
                                   const double *samp, const double *x, const double *y,
                                   const double *z, double prec, const LocusOut &o,
                                   cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  size_t smemBytes;
  bool stage;
  LaunchShape sh = shapeFor(S, n, smCount, smemBytes, stage, 3);
  if (stage) {
    if (smemBytes > 48 * 1024)
      cudaFuncSetAttribute(lsProximateLocusKernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)smemBytes);
    lsProximateLocusKernel<true><<<sh.blocks, sh.threads, smemBytes, stream>>>(S, n, line, samp, x, y,
                                                                               z, prec, o);
  } else {
    lsProximateLocusKernel<false><<<sh.blocks, sh.threads, 0, stream>>>(S, n, line, samp, x, y, z,
                                                                        prec, o);
  }
  countLaunch();
  return cudaGetLastError();
}

}  // namespace ub200
