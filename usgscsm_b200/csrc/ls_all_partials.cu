// ls_all_partials.cu -- line-scan computeAllSensorPartials in ONE kernel.
//
// Replaces csm::RasterGM::computeAllSensorPartials (csm/RasterGM.cpp:52-140, reached through
// UsgsAstroLsSensorModel::computeAllSensorPartials LS:1413-1455): groundToImage of the ground
// point, then one more groundToImage per requested parameter with that parameter moved by m_gsd
// (computeSensorPartials LS:1373-1396), forward differences.
//
// What the fusion shares.  Every one of those searches starts from the SAME projective seed
// (it does not depend on the parameters) and evaluates its residual first at seed + 0 and at
// seed + 0.1 lines (LS:810-813) -- the same two image times for all of them.  The interpolated
// ephemeris position / velocity, the quaternion, the camera-to-body matrix and the in-track /
// cross-track / radial frame at those two times are therefore computed ONCE per point and only
// the cheap parameter-dependent tail (bias in that frame, Euler correction, focal length,
// distortion, detector line) runs per search; each search then continues on its own (one more
// evaluation at its own time, typically).  17 searches cost 2 + 17 interpolations instead of 51.
// Same arithmetic on the same values as the separate searches, hence the same bits.
#include <cstdlib>
#include <type_traits>

#include "kernels.h"
#include "ls_common.cuh"
#include "work_queue.cuh"

namespace ub200 {

namespace {

constexpr int kThreads = 256;
constexpr int kMaxIndices = 16;

struct IndexList {
  int count;
  int index[kMaxIndices];
};

// What depends on the image time only.
struct NominalView {
  double pos[3], vel[3];
  double ecf[9];  // in-track / cross-track / radial frame (only when some search needs it)
  double m[9];    // camera to body
};

// The pieces below are NOT inlined: the kernel calls each from two places (the shared first two
// evaluations and every search's own continuation), and with everything inlined its code grew to
// several times the instruction cache (ncu: instruction-cache hit rate 58 %, "no instruction" 8
// stall cycles per issue).  A call costs a few cycles against bodies of 200 - 500 instructions.
template <class TAB, class SPEC>
__device__ __noinline__ void nominalView(const LsState &S, const TAB &T, double time, bool needFrame,
                                         NominalView &v) {
  LagrangeWeights w;
  lagrangeWeights<SPEC::kColdOutOfLine>(S.positions, S.lag, time, w);
  tableApply<3>(T, T.positions, w, v.pos);
  if (needFrame) {
    tableApply<3>(T, T.velocities, w, v.vel);
    icrFrame(v.pos, v.vel, v.ecf);
  }
  cameraToBodyMatrix<TAB, SPEC>(S, T, time, v.m);
}

// Parameter j of the search that moves parameter idx by delta: getValue(j, adj) with
// adj[idx] = delta (LS:1383-1389); idx = -1: the model as it is.
UB_DEV double movedParam(const LsState &S, int idx, double delta, int j) {
  return S.params[j] + (j == idx ? delta : 0.0);
}
UB_DEV void movedFlags(const LsState &S, int idx, double delta, bool &posAdj, bool &attAdj) {
  posAdj = attAdj = false;
#pragma unroll
  for (int j = 0; j < 6; j++) posAdj = posAdj || movedParam(S, idx, delta, j) != 0;
#pragma unroll
  for (int j = 6; j < 15; j++) attAdj = attAdj || movedParam(S, idx, delta, j) != 0;
}

// computeDetectorView LS:2603-2675 from a nominal view with the moved parameters, then
// calcDetectorLineErr's tail LS:3221-3229 (distortion at the default precision, detector line).
// When none of parameters 0..5 is non-zero the reference's adjusted position is the nominal one
// exactly (it adds products with exact zeros: LS:2567-2583), so the frame is not needed; with
// parameters 6..14 all zero the attitude correction is the identity.
template <int DIST>
__device__ __noinline__ double movedResidual(const LsState &S, const HoistedScaleDivision &div,
                                             const NominalView &v, double time, int idx, double delta,
                                             double gx, double gy, double gz, double &focalX,
                                             double &focalY) {
  bool posAdj, attAdj;
  movedFlags(S, idx, delta, posAdj, attAdj);
  auto pa = [&](int j) { return movedParam(S, idx, delta, j); };
  double pos[3], vel[3];
  if (posAdj) {
    applyPositionAdjustment(S, v.ecf, v.pos, v.vel, time, pa, pos, vel, div);
  } else {
    pos[0] = v.pos[0];
    pos[1] = v.pos[1];
    pos[2] = v.pos[2];
  }
  double bodyLook[3] = {gx - pos[0], gy - pos[1], gz - pos[2]};
  double cameraLook[3];
  matTVec(v.m, bodyLook, cameraLook);
  double adjustedLook[3];
  if (attAdj) {
    double attCorr[9];
    attitudeCorrectionWith(S, time, pa, attCorr, div);
    matTVec(attCorr, cameraLook, adjustedLook);
  } else {
    adjustedLook[0] = cameraLook[0];
    adjustedLook[1] = cameraLook[1];
    adjustedLook[2] = cameraLook[2];
  }
  double lookScale = (S.fp.focalLength + pa(15)) / adjustedLook[2];
  focalX = adjustedLook[0] * lookScale;
  focalY = adjustedLook[1] * lookScale;
  const FocalPlane &fp = S.fp;
  double dx, dy;
  uint32_t st = 0;
  applyDistortion<DIST>(focalX, focalY, dx, dy, fp, 1.0E-6, st);
  double detectorLine =
      fp.iTransL[0] + fp.iTransL[1] * dx + fp.iTransL[2] * dy + fp.detectorLineOrigin - fp.startingDetectorLine;
  return divideBy(detectorLine, fp.lineSumming);
}

template <bool STAGE, int DIST, int MINB>
__global__ void __launch_bounds__(kThreads, MINB)
    lsAllSensorPartialsKernel(const __grid_constant__ LsState S, const __grid_constant__ IndexList L,
                              long n, const double *__restrict__ x, const double *__restrict__ y,
                              const double *__restrict__ z, const double *__restrict__ imgLine,
                              const double *__restrict__ imgSamp, double prec,
                              double *__restrict__ partials, double *__restrict__ achievedOut,
                              uint32_t *__restrict__ statusOut, unsigned long long *__restrict__ next,
                              int grab) {
  extern __shared__ __align__(16) double smem[];
  using SPEC = typename std::conditional<(DIST >= 0), CommonSpec<DIST, -1>, SearchSpec>::type;
  using TAB = typename std::conditional<STAGE, StagedTables, OrientationTables>::type;
  TAB T{S.positions.values, S.velocities.values, S.quaternions.values};
  if (STAGE) {
    StagedTables staged = stageOrientationTables(S, true, smem);
    T.positions = staged.positions;
    T.velocities = staged.velocities;
    T.quaternions = staged.quaternions;
  }
  const FocalPlane &fp = S.fp;
  const double delta = S.gsd;
  const HoistedScaleDivision div(S);
  const bool haveImage = imgLine != nullptr;
  const uint32_t kErr = USGSCSM_B200_ST_ERROR_ALGORITHM | USGSCSM_B200_ST_ERROR_UNKNOWN |
                        USGSCSM_B200_ST_ERROR_OTHER;
  const uint32_t kKeep = kErr | USGSCSM_B200_ST_NO_INTERSECTION | USGSCSM_B200_ST_PRECISION_NOT_MET;
  const int K = L.count;
  // slot s of the per-search arrays: s < K moves parameter L.index[s]; s == K is the model as it is
  const int nSearch = haveImage ? K : K + 1;

  forEachPointDynamic(n, next, grab, [&](long i) {
    const double gx = __ldcs(x + i), gy = __ldcs(y + i), gz = __ldcs(z + i);
    double approxLine, approxSamp;
    projectiveSeed<SPEC>(S, gx, gy, gz, approxLine, approxSamp);
    bool anyFrame = false;
    for (int s = 0; s < nSearch; s++) {
      bool p, a;
      movedFlags(S, s < K ? L.index[s] : -1, delta, p, a);
      anyFrame = anyFrame || p;
    }

    // the first two evaluations of every search: residuals at seed + 0 and seed + 0.1 lines
    double err0[kMaxIndices + 1], err1[kMaxIndices + 1], fx1[kMaxIndices + 1], fy1[kMaxIndices + 1];
    // (the view is a stack variable: keeping it in shared memory instead cost a resident block per
    // SM and measured 15 % slower, profiles/r2_ab_partials.txt)
    NominalView v;
#pragma unroll 1
    for (int e = 0; e < 2; e++) {
      const double time = lsImageTime<SPEC>(S, approxLine + (e ? 0.1 : 0.0));
      nominalView<TAB, SPEC>(S, T, time, anyFrame, v);
#pragma unroll 1
      for (int s = 0; s < nSearch; s++) {
        double fx, fy;
        const double err =
            movedResidual<SPEC::kDist>(S, div, v, time, s < K ? L.index[s] : -1, delta, gx, gy, gz, fx, fy);
        if (e == 0) {
          err0[s] = err;
        } else {
          err1[s] = err;
          fx1[s] = fx;
          fy1[s] = fy;
        }
      }
    }

    // every search from its third evaluation on (LS:817-833) and its final projection
    // (LS:838-890); q = -1: the image point itself, unless the caller gave it
    double l0 = 0.0, s0 = 0.0, am = 0.0;
    uint32_t acc = 0;
    bool finite = true;
    if (haveImage) {
      l0 = __ldcs(imgLine + i);
      s0 = __ldcs(imgSamp + i);
    }
    double *out = partials + (size_t)i * K * 2;
#pragma unroll 1
    for (int q = haveImage ? 0 : -1; q < K; q++) {
      const int s = q < 0 ? K : q;
      const int idx = q < 0 ? -1 : L.index[q];
      bool posAdj, attAdj;
      movedFlags(S, idx, delta, posAdj, attAdj);
      double lineErr0 = err0[s], lineErr1 = err1[s], focalX = fx1[s], focalY = fy1[s];
      uint32_t st = 0;
      int count = 0;
      double t0 = 0.0, t1 = 0.1;
      for (;;) {
        if (!(fabs(lineErr1) > prec)) break;
        if (!(++count < 15)) {
          st |= USGSCSM_B200_ST_ITERATION_CAP;
          break;
        }
        if (lineErr1 == lineErr0) {
          st |= USGSCSM_B200_ST_SEARCH_STALL;
          break;
        }
        double t2 = t1 - lineErr1 * (t1 - t0) / (lineErr1 - lineErr0);
        t0 = t1;
        lineErr0 = lineErr1;
        t1 = t2;
        const double time = lsImageTime<SPEC>(S, approxLine + t1);
        nominalView<TAB, SPEC>(S, T, time, posAdj, v);
        lineErr1 = movedResidual<SPEC::kDist>(S, div, v, time, idx, delta, gx, gy, gz, focalX, focalY);
      }
      const double lineAt = approxLine + t1;
      double dx, dy;
      uint32_t distStatus = 0;
      applyDistortion<SPEC::kDist>(focalX, focalY, dx, dy, fp, prec, distStatus);
      st |= distStatus;
      double detectorLine = fp.iTransL[0] + fp.iTransL[1] * dx + fp.iTransL[2] * dy;
      double detectorSample = fp.iTransS[0] + fp.iTransS[1] * dx + fp.iTransS[2] * dy;
      double finalUpdate =
          divideBy(detectorLine + fp.detectorLineOrigin - fp.startingDetectorLine, fp.lineSumming);
      const double l1 = lineAt + finalUpdate;
      const double s1 =
          divideBy(detectorSample + fp.detectorSampleOrigin - fp.startingDetectorSample, fp.sampleSumming);
      if ((prec > 0.0) && (fabs(finalUpdate) > prec)) st |= USGSCSM_B200_ST_PRECISION_NOT_MET;
      acc |= st & kKeep;
      if (q < 0) {
        l0 = l1;
        s0 = s1;
      } else {
        if (finalUpdate > am) am = finalUpdate;  // achieved precision = finalUpdate (LS:874-876)
        const double pl = (l1 - l0) / delta;
        const double ps = (s1 - s0) / delta;
        out[2 * q] = pl;
        out[2 * q + 1] = ps;
        finite = finite && isfinite(pl) && isfinite(ps);
      }
    }
    if (!finite) acc |= USGSCSM_B200_ST_NONFINITE;
    if (acc & kErr) {  // a scalar call would have thrown: the point has no list
      acc = (acc & kErr) | USGSCSM_B200_ST_NONFINITE;
      for (int q = 0; q < 2 * K; q++) out[q] = nan("");
      am = nan("");
    }
    if (statusOut) statusOut[i] = acc;
    if (achievedOut) achievedOut[i] = am;
  });
}

int envInt(const char *name, int def) {
  const char *v = getenv(name);
  return v && *v ? atoi(v) : def;
}

}  // namespace

// indices: the parameters to move, in output order (n_indices <= 16).  False when the fused
// kernel does not apply (USGSCSM_B200_LS_FUSED_PARTIALS=0) and the caller should run the
// per-parameter launches.
bool launchLsAllSensorPartials(const LsState &S, int smCount, int nIndices, const int *indices, long n,
                               const double *x, const double *y, const double *z,
                               const double *imgLine, const double *imgSamp, double prec,
                               double *partials, double *achieved, uint32_t *status,
                               cudaStream_t stream, cudaError_t &err) {
  // Off by default: the per-parameter searches (each a launch of the specialised search kernel
  // that writes its forward difference itself) measure 7.2 G searches/s against this kernel's 5.1
  // -- it carries 1 KB of per-thread state (profiles/r2_ab_partials.txt, r2_ncu_all_partials.txt).
  static const int fused = envInt("USGSCSM_B200_LS_FUSED_PARTIALS", 0);
  err = cudaSuccess;
  if (!fused || nIndices < 1 || nIndices > kMaxIndices) return false;
  if (n <= 0) return true;
  IndexList L{};
  L.count = nIndices;
  for (int k = 0; k < nIndices; k++) L.index[k] = indices ? indices[k] : k;
  const size_t bytes = orientationTableBytes(S, true);
  const bool stage = bytes <= kMaxStagedTableBytes;
  const size_t smemBytes = stage ? bytes : 0;
  long blocksNeeded = (n + kThreads - 1) / kThreads;
  // register budget as resident blocks per SM (USGSCSM_B200_LS_PARTIALS_MINB: 2, 3 or 4)
  static const int minb = envInt("USGSCSM_B200_LS_PARTIALS_MINB", 4);
  const int mb = minb < 2 ? 2 : (minb > 4 ? 4 : minb);
  long persistent = (long)smCount * mb;
  const int blocks = (int)(blocksNeeded < persistent ? blocksNeeded : persistent);
  WorkCounter wc;
  err = wc.acquire(stream, n);
  if (err != cudaSuccess) return true;
  static const int specialise = envInt("USGSCSM_B200_LS_SPECIALISE", 1);
  const bool common = specialise == 1 && S.numIntTimes == 1 && S.useApproxInitTrans;
#define UB_LAUNCH_AP_M(STAGE, DIST, MB)                                                             \
  lsAllSensorPartialsKernel<STAGE, DIST, MB><<<blocks, kThreads, smemBytes, stream>>>(              \
      S, L, n, x, y, z, imgLine, imgSamp, prec, partials, achieved, status, wc.ptr, 64)
#define UB_LAUNCH_AP(STAGE, DIST)                     \
  do {                                                \
    if (mb <= 2) UB_LAUNCH_AP_M(STAGE, DIST, 2);      \
    else if (mb == 3) UB_LAUNCH_AP_M(STAGE, DIST, 3); \
    else UB_LAUNCH_AP_M(STAGE, DIST, 4);              \
  } while (0)
#define UB_LAUNCH_AP_DIST(STAGE)                                          \
  do {                                                                    \
    if (!common) UB_LAUNCH_AP(STAGE, -1);                                 \
    else if (S.fp.distortionType == 4) UB_LAUNCH_AP(STAGE, 4);            \
    else if (S.fp.distortionType == 0) UB_LAUNCH_AP(STAGE, 0);            \
    else if (S.fp.distortionType == 1 && transverseCoeffsPlain(S.fp)) UB_LAUNCH_AP(STAGE, 1);            \
    else UB_LAUNCH_AP(STAGE, -1);                                         \
  } while (0)
  if (stage)
    UB_LAUNCH_AP_DIST(true);
  else
    UB_LAUNCH_AP_DIST(false);
#undef UB_LAUNCH_AP_DIST
#undef UB_LAUNCH_AP
#undef UB_LAUNCH_AP_M
  countLaunch();
  err = wc.release();
  return true;
}

}  // namespace ub200
