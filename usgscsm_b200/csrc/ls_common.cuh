// ls_common.cuh -- exterior orientation and ray geometry shared by the line-scan and
// push-frame kernels (the reference's two models are near copies of each other:
// src/UsgsAstroLsSensorModel.cpp and src/UsgsAstroPushFrameSensorModel.cpp).
#ifndef USGSCSM_B200_LS_COMMON_CUH_
#define USGSCSM_B200_LS_COMMON_CUH_

#include "device_math.cuh"

namespace ub200 {

// Where the kernels read the ephemeris / attitude tables from: the device arrays of the
// packed state (rows of 3 / 4 doubles), or a copy the block staged into shared memory as
// rows of 4 doubles on 16-byte boundaries (128-bit loads).
struct OrientationTables {
  static constexpr bool kPadded = false;
  const double *positions;
  const double *velocities;
  const double *quaternions;
};
struct StagedTables {
  static constexpr bool kPadded = true;
  const double *positions;
  const double *velocities;
  const double *quaternions;
};

template <int VLEN, class TAB>
UB_DEV void tableApply(const TAB &, const double *table, const LagrangeWeights &w,
                       double out[VLEN]) {
  if (TAB::kPadded)
    lagrangeApplyPadded<VLEN>(table, w, out);
  else
    lagrangeApply<VLEN>(table, w, out);
}

// Bytes of shared memory needed to stage the tables of S (velocities only when the
// adjusted-position path can run).
__host__ __device__ inline size_t orientationTableBytes(const LsState &S, bool needVelocity) {
  size_t n = (size_t)S.positions.numNodes * 4 + (size_t)S.quaternions.numNodes * 4;
  if (needVelocity) n += (size_t)S.velocities.numNodes * 4;
  return n * sizeof(double);
}

// Cooperative copy of the tables into `smem` (16-byte aligned); returns the staged pointers.
// All threads of the block must call it; ends with __syncthreads().
__device__ inline StagedTables stageOrientationTables(const LsState &S, bool needVelocity,
                                                      double *smem) {
  StagedTables t;
  const int np = S.positions.numNodes;
  const int nq = S.quaternions.numNodes * 4;
  double *sp = smem;
  double *sq = sp + 4 * np;
  double *sv = sq + nq;
  for (int i = threadIdx.x; i < 4 * np; i += blockDim.x) {
    int node = i >> 2, j = i & 3;
    sp[i] = j < 3 ? __ldg(S.positions.values + 3 * node + j) : 0.0;
  }
  for (int i = threadIdx.x; i < nq; i += blockDim.x) sq[i] = __ldg(S.quaternions.values + i);
  if (needVelocity) {
    const int nv = S.velocities.numNodes;
    for (int i = threadIdx.x; i < 4 * nv; i += blockDim.x) {
      int node = i >> 2, j = i & 3;
      sv[i] = j < 3 ? __ldg(S.velocities.values + 3 * node + j) : 0.0;
    }
  }
  __syncthreads();
  t.positions = sp;
  t.quaternions = sq;
  t.velocities = sv;  // only read when needVelocity (S.hasAdj)
  return t;
}

// What a kernel may know about the model at compile time.  The generic spec reads everything
// from the packed state; CommonSpec<DIST> is the common case the launcher can verify on the
// host -- no parameter adjustments, a single line-rate segment, projective seed available,
// distortion model DIST -- so that every other branch drops out of the instruction stream.
struct GenericSpec {
  static constexpr bool kCommon = false;
  static constexpr int kDist = -1;
  static constexpr bool kColdOutOfLine = false;  // see lagrangeWeights
  static constexpr int kAdj = -1;                // read hasAdj / hasAttAdj from the state
};
struct SearchSpec : GenericSpec {  // the line-scan search kernel, any model
  static constexpr bool kColdOutOfLine = true;
};
// ADJ: bit 0 = some of the position / velocity parameters (0..5) are set, bit 1 = some of the
// attitude parameters (6..14) are; 3 = the state of a model after bundle adjustment.
template <int DIST, int ADJ = 0>
struct CommonSpec {
  static constexpr bool kCommon = true;
  static constexpr int kDist = DIST;
  static constexpr bool kColdOutOfLine = true;
  static constexpr int kAdj = ADJ;
};
template <class SPEC>
UB_DEV bool specHasAdj(const LsState &S) { return SPEC::kAdj < 0 ? S.hasAdj != 0 : (SPEC::kAdj & 1) != 0; }
template <class SPEC>
UB_DEV bool specHasAttAdj(const LsState &S) {
  return SPEC::kAdj < 0 ? S.hasAttAdj != 0 : (SPEC::kAdj & 2) != 0;
}

// getValue(i, adj) with adj == _no_adjustment (LS:757-772 passes 16 zeros): the host packs
// m_currentParameterValue[i] + 0.0.
UB_DEV double param(const LsState &S, int i) { return S.params[i]; }

// LS::getImageTime LS:1543-1568: std::upper_bound over m_intTimeLines, step back one.
template <class SPEC = GenericSpec>
UB_DEV double lsImageTime(const LsState &S, double lineFull) {
  if (SPEC::kCommon || S.numIntTimes == 1) {
    return S.intStart0 + S.intTime0 * (lineFull - S.intLine0 + 0.5);
  }
  // std::upper_bound: first element with lineFull < element
  int first = 0;
  int len = S.numIntTimes;
  while (len > 0) {
    int half = len >> 1;
    int mid = first + half;
    if (lineFull < S.intTimeLines[mid]) {
      len = half;
    } else {
      first = mid + 1;
      len = len - half - 1;
    }
  }
  int idx = first;
  if (idx != 0) --idx;
  return S.intTimeStartTimes[idx] + S.intTimes[idx] * (lineFull - S.intTimeLines[idx] + 0.5);
}

// The in-track / cross-track / radial frame of LS::getAdjSensorPosVel LS:2530-2565 (PF:1994-2020)
// from the nominal position and velocity: columns of ecfFromIcr.
UB_DEV void icrFrame(const double sensPosNom[3], const double sensVelNom[3], double ecfFromIcr[9]) {
  double radialUnitVec[3];
  double radMag = sqrt(sensPosNom[0] * sensPosNom[0] + sensPosNom[1] * sensPosNom[1] +
                       sensPosNom[2] * sensPosNom[2]);
  double rRadMag = refinedReciprocal(radMag);
  for (int i = 0; i < 3; i++) radialUnitVec[i] = divideBy(sensPosNom[i], radMag, rRadMag);
  double crossTrackUnitVec[3];
  crossTrackUnitVec[0] = sensPosNom[1] * sensVelNom[2] - sensPosNom[2] * sensVelNom[1];
  crossTrackUnitVec[1] = sensPosNom[2] * sensVelNom[0] - sensPosNom[0] * sensVelNom[2];
  crossTrackUnitVec[2] = sensPosNom[0] * sensVelNom[1] - sensPosNom[1] * sensVelNom[0];
  double crossMag = sqrt(crossTrackUnitVec[0] * crossTrackUnitVec[0] +
                         crossTrackUnitVec[1] * crossTrackUnitVec[1] +
                         crossTrackUnitVec[2] * crossTrackUnitVec[2]);
  double rCrossMag = refinedReciprocal(crossMag);
  for (int i = 0; i < 3; i++)
    crossTrackUnitVec[i] = divideBy(crossTrackUnitVec[i], crossMag, rCrossMag);
  double inTrackUnitVec[3];
  inTrackUnitVec[0] =
      crossTrackUnitVec[1] * radialUnitVec[2] - crossTrackUnitVec[2] * radialUnitVec[1];
  inTrackUnitVec[1] =
      crossTrackUnitVec[2] * radialUnitVec[0] - crossTrackUnitVec[0] * radialUnitVec[2];
  inTrackUnitVec[2] =
      crossTrackUnitVec[0] * radialUnitVec[1] - crossTrackUnitVec[1] * radialUnitVec[0];
  ecfFromIcr[0] = inTrackUnitVec[0];
  ecfFromIcr[1] = crossTrackUnitVec[0];
  ecfFromIcr[2] = radialUnitVec[0];
  ecfFromIcr[3] = inTrackUnitVec[1];
  ecfFromIcr[4] = crossTrackUnitVec[1];
  ecfFromIcr[5] = radialUnitVec[1];
  ecfFromIcr[6] = inTrackUnitVec[2];
  ecfFromIcr[7] = crossTrackUnitVec[2];
  ecfFromIcr[8] = radialUnitVec[2];
}

// Division by the model's derived scales (m_halfTime, m_flyingHeight, m_halfSwath) in the
// adjustment formulas: as written, or with reciprocals the caller hoisted out of its point loop
// (divideBy returns the IEEE quotient either way).
struct ScaleDivision {
  UB_DEV ScaleDivision() {}
  UB_DEV explicit ScaleDivision(const LsState &) {}
  UB_DEV double byHalfTime(const LsState &S, double a) const { return a / S.halfTime; }
  UB_DEV double byFlyingHeight(const LsState &S, double a) const { return a / S.flyingHeight; }
  UB_DEV double byHalfSwath(const LsState &S, double a) const { return a / S.halfSwath; }
};
struct HoistedScaleDivision {
  double rHalfTime, rFlyingHeight, rHalfSwath;
  UB_DEV explicit HoistedScaleDivision(const LsState &S)
      : rHalfTime(refinedReciprocal(S.halfTime)), rFlyingHeight(refinedReciprocal(S.flyingHeight)),
        rHalfSwath(refinedReciprocal(S.halfSwath)) {}
  UB_DEV double byHalfTime(const LsState &S, double a) const { return divideBy(a, S.halfTime, rHalfTime); }
  UB_DEV double byFlyingHeight(const LsState &S, double a) const { return divideBy(a, S.flyingHeight, rFlyingHeight); }
  UB_DEV double byHalfSwath(const LsState &S, double a) const { return divideBy(a, S.halfSwath, rHalfSwath); }
};

// The bias / rate adjustments of LS:2567-2583 applied in that frame; pa(i) = getValue(i, adj).
template <class PA, class DIV = ScaleDivision>
UB_DEV void applyPositionAdjustment(const LsState &S, const double ecfFromIcr[9],
                                    const double sensPosNom[3], const double sensVelNom[3],
                                    double time, PA &&pa, double pos[3], double vel[3],
                                    const DIV &div = DIV()) {
  double aTime = time - S.positions.t0;
  double dvi = div.byHalfTime(S, pa(3));
  double dvc = div.byHalfTime(S, pa(4));
  double dvr = div.byHalfTime(S, pa(5));
  vel[0] = sensVelNom[0] + ecfFromIcr[0] * dvi + ecfFromIcr[1] * dvc + ecfFromIcr[2] * dvr;
  vel[1] = sensVelNom[1] + ecfFromIcr[3] * dvi + ecfFromIcr[4] * dvc + ecfFromIcr[5] * dvr;
  vel[2] = sensVelNom[2] + ecfFromIcr[6] * dvi + ecfFromIcr[7] * dvc + ecfFromIcr[8] * dvr;
  double di = pa(0) + dvi * aTime;
  double dc = pa(1) + dvc * aTime;
  double dr = pa(2) + dvr * aTime;
  pos[0] = sensPosNom[0] + ecfFromIcr[0] * di + ecfFromIcr[1] * dc + ecfFromIcr[2] * dr;
  pos[1] = sensPosNom[1] + ecfFromIcr[3] * di + ecfFromIcr[4] * dc + ecfFromIcr[5] * dr;
  pos[2] = sensPosNom[2] + ecfFromIcr[6] * di + ecfFromIcr[7] * dc + ecfFromIcr[8] * dr;
}

// LS::getAdjSensorPosVel LS:2484-2589 / PF:1955-2034.
//   ALWAYS_VEL = false: line-scan semantics (calc_vel == false at every hot-path call site:
//                position only unless some parameter is non-zero);
//   ALWAYS_VEL = true:  push-frame semantics (no shortcut).
template <bool ALWAYS_VEL, class TAB, class SPEC = GenericSpec, class DIV = ScaleDivision>
UB_DEV void adjSensorPosVel(const LsState &S, const TAB &T, double time,
                            double pos[3], double vel[3], const DIV &div = DIV()) {
  LagrangeWeights w;
  lagrangeWeights<SPEC::kColdOutOfLine>(S.positions, S.lag, time, w);
  double sensPosNom[3];
  tableApply<3>(T, T.positions, w, sensPosNom);
  if (!ALWAYS_VEL && !specHasAdj<SPEC>(S)) {
    pos[0] = sensPosNom[0];
    pos[1] = sensPosNom[1];
    pos[2] = sensPosNom[2];
    vel[0] = 0.0;
    vel[1] = 0.0;
    vel[2] = 0.0;
    return;
  }
  // Same grid, same time, same order => the same weights (the reference recomputes them).
  double sensVelNom[3];
  tableApply<3>(T, T.velocities, w, sensVelNom);
  double ecfFromIcr[9];
  icrFrame(sensPosNom, sensVelNom, ecfFromIcr);
  applyPositionAdjustment(S, ecfFromIcr, sensPosNom, sensVelNom, time,
                          [&](int i) { return param(S, i); }, pos, vel, div);
}

// LS::getQuaternions LS:2140-2153 + calculateRotationMatrixFromQuaternions U:78-95.
template <class TAB, class SPEC = GenericSpec>
UB_DEV void cameraToBodyMatrix(const LsState &S, const TAB &T, double time, double m[9]) {
  double q[4];
  LagrangeWeights w;
  lagrangeWeights<SPEC::kColdOutOfLine>(S.quaternions, S.lag, time, w);
  tableApply<4>(T, T.quaternions, w, q);
  rotationFromQuaternion(q, m);
}

// LS::calculateAttitudeCorrection LS:2162-2187.  Only called when S.hasAttAdj: with
// parameters 6..14 all zero the Euler angles are exactly 0 and the matrix is the identity
// up to the sign of its zeros, so applying it returns its argument unchanged.
template <class PA, class DIV = ScaleDivision>
UB_DEV void attitudeCorrectionWith(const LsState &S, double time, PA &&pa, double attCorr[9],
                                   const DIV &div = DIV()) {
  double aTime = time - S.quaternions.t0;
  double euler[3];
  double nTime = div.byHalfTime(S, aTime);
  double nTime2 = nTime * nTime;
  euler[0] = div.byFlyingHeight(S, pa(6) + pa(9) * nTime + pa(12) * nTime2);
  euler[1] = div.byFlyingHeight(S, pa(7) + pa(10) * nTime + pa(13) * nTime2);
  euler[2] = div.byHalfSwath(S, pa(8) + pa(11) * nTime + pa(14) * nTime2);
  rotationFromEuler(euler, attCorr);
}
template <class DIV = ScaleDivision>
UB_DEV void attitudeCorrection(const LsState &S, double time, double attCorr[9], const DIV &div = DIV()) {
  attitudeCorrectionWith(S, time, [&](int i) { return param(S, i); }, attCorr, div);
}

// LS::computeDetectorView LS:2603-2675 (PF:2038-2110 is the same with the push-frame
// position routine): ground point -> undistorted focal-plane coordinates at `time`.
template <bool ALWAYS_VEL, class TAB, class SPEC = GenericSpec, class DIV = ScaleDivision>
UB_DEV void detectorView(const LsState &S, const TAB &T, double time, double gx,
                         double gy, double gz, double &focalX, double &focalY,
                         const DIV &div = DIV()) {
  double pos[3], vel[3];
  adjSensorPosVel<ALWAYS_VEL, TAB, SPEC, DIV>(S, T, time, pos, vel, div);
  double bodyLook[3] = {gx - pos[0], gy - pos[1], gz - pos[2]};
  double bodyToCamera[9];
  cameraToBodyMatrix<TAB, SPEC>(S, T, time, bodyToCamera);
  double cameraLook[3];
  matTVec(bodyToCamera, bodyLook, cameraLook);
  double adjustedLook[3];
  if (specHasAttAdj<SPEC>(S)) {
    double attCorr[9];
    attitudeCorrection(S, time, attCorr, div);
    matTVec(attCorr, cameraLook, adjustedLook);
  } else {
    adjustedLook[0] = cameraLook[0];
    adjustedLook[1] = cameraLook[1];
    adjustedLook[2] = cameraLook[2];
  }
  double lookScale = (S.fp.focalLength + param(S, 15)) / adjustedLook[2];
  focalX = adjustedLook[0] * lookScale;
  focalY = adjustedLook[1] * lookScale;
}

// Camera-frame look vector -> body-fixed look (second half of losToEcf, LS:2268-2300).
template <class TAB>
UB_DEV void cameraLookToBody(const LsState &S, const TAB &T, double time,
                             const double cameraLook[3], double bodyLook[3]) {
  double corrected[3];
  if (S.hasAttAdj) {
    double attCorr[9];
    attitudeCorrection(S, time, attCorr);
    matVec(attCorr, cameraLook, corrected);
  } else {
    corrected[0] = cameraLook[0];
    corrected[1] = cameraLook[1];
    corrected[2] = cameraLook[2];
  }
  double cameraToBody[9];
  cameraToBodyMatrix(S, T, time, cameraToBody);
  matVec(cameraToBody, corrected, bodyLook);
}

// LS::losEllipsoidIntersect LS:2398-2465 (PF:1883-1950 identical): root with the smaller
// magnitude; negative discriminant clamped with NO_INTERSECTION; achieved precision from
// computeEllipsoidElevation.
UB_DEV void lsEllipsoidIntersect(const Ellipsoid &ell, double height, const double c[3],
                                 const double l[3], double desired_precision, double &x,
                                 double &y, double &z, double &achieved, uint32_t &status) {
  double ap = ell.majorAxis + height;
  double bp = ell.minorAxis + height;
  // sphere: ap and bp are the same double, and x / x == 1 for any finite non-zero x
  double ap2 = ap * ap;
  double k = (ell.majorAxis == ell.minorAxis && ap2 > 0.0 && ap2 < CUDART_INF) ? 1.0 : ap2 / (bp * bp);
  double at = l[0] * l[0] + l[1] * l[1] + k * l[2] * l[2];
  double bt = 2.0 * (l[0] * c[0] + l[1] * c[1] + k * l[2] * c[2]);
  double ct = c[0] * c[0] + c[1] * c[1] + k * c[2] * c[2] - ap * ap;
  double quadTerm = bt * bt - 4.0 * at * ct;
  if (0.0 > quadTerm) {
    quadTerm = 0.0;
    status |= USGSCSM_B200_ST_NO_INTERSECTION;
  }
  double sTerm = sqrt(quadTerm);
  double scale = (-bt - sTerm);
  double scale1 = (-bt + sTerm);
  if (fabs(scale1) < fabs(scale)) scale = scale1;
  scale /= (2.0 * at);
  x = c[0] + scale * l[0];
  y = c[1] + scale * l[1];
  z = c[2] + scale * l[2];
  double h = ellipsoidElevation(x, y, z, ell, desired_precision);
  achieved = fabs(height - h);
}

// LS::computeProjectiveApproximation LS:2688-2736
template <class SPEC = GenericSpec>
UB_DEV void projectiveSeed(const LsState &S, double x, double y, double z, double &line,
                           double &samp) {
  if (SPEC::kCommon || S.useApproxInitTrans) {
    const double *u = S.projCoeffs;
    double line_den = 1 + u[4] * x + u[5] * y + u[6] * z;
    double samp_den = 1 + u[11] * x + u[12] * y + u[13] * z;
    if (line_den == 0.0 || isnan(line_den) || isinf(line_den) || samp_den == 0.0 ||
        isnan(samp_den) || isinf(samp_den)) {
      line = S.nLines / 2.0;
      samp = S.nSamples / 2.0;
      return;
    }
    line = (u[0] + u[1] * x + u[2] * y + u[3] * z) / line_den;
    samp = (u[7] + u[8] * x + u[9] * y + u[10] * z) / samp_den;
    double numRows = S.nLines;
    double numCols = S.nSamples;
    if (line < 0.0) line = 0.0;
    if (line > numRows) line = numRows;
    if (samp < 0.0) samp = 0.0;
    if (samp > numCols) samp = numCols;
  } else {
    line = S.nLines / 2.0;
    samp = S.nSamples / 2.0;
  }
}

UB_DEV uint32_t imageRangeStatus(double line, double samp, int nLines, int nSamples) {
  uint32_t st = 0;
  if (!(isfinite(line) && isfinite(samp))) st |= USGSCSM_B200_ST_NONFINITE;
  // getValidImageRange LS:1862-1869: (0,0) .. (nLines, nSamples), inclusive
  if (!(line >= 0.0 && line <= (double)nLines && samp >= 0.0 && samp <= (double)nSamples))
    st |= USGSCSM_B200_ST_OUT_OF_IMAGE;
  return st;
}

// imageToProximateImagingLocus LS:1175-1235 / PF:995-1055: three imageToGround calls (at the
// ground point's height, one gsd below it, and at the height of the point of the ray closest to
// the ground point) and two ellipsoid elevations.  `i2g(height, xyz, achieved, status)` is the
// model's imageToGround for the pixel.  Only the last call receives the caller's warning
// list; a csm::Error thrown by any of them escapes (outputs NaN).
template <class I2G>
UB_DEV void proximateLocusFromImageToGround(const Ellipsoid &ell, double gsd, double x, double y,
                                            double z, double desiredPrecision, I2G &&i2g,
                                            double pos[3], double dir[3], double &achieved,
                                            uint32_t &status) {
  const uint32_t kErr = USGSCSM_B200_ST_ERROR_ALGORITHM | USGSCSM_B200_ST_ERROR_UNKNOWN |
                        USGSCSM_B200_ST_ERROR_OTHER;
  double height = ellipsoidElevation(x, y, z, ell, desiredPrecision);
  double gp1[3], gp2[3], a;
  uint32_t st1 = 0, st2 = 0, st3 = 0;
  i2g(height, gp1, a, st1);
  double dx1 = x - gp1[0];
  double dy1 = y - gp1[1];
  double dz1 = z - gp1[2];
  i2g(height - gsd, gp2, a, st2);
  double dx2 = gp2[0] - gp1[0];
  double dy2 = gp2[1] - gp1[1];
  double dz2 = gp2[2] - gp1[2];
  double mag2 = sqrt(dx2 * dx2 + dy2 * dy2 + dz2 * dz2);
  dx2 /= mag2;
  dy2 /= mag2;
  dz2 /= mag2;
  double scale = dx1 * dx2 + dy1 * dy2 + dz1 * dz2;
  gp2[0] = gp1[0] + scale * dx2;
  gp2[1] = gp1[1] + scale * dy2;
  gp2[2] = gp1[2] + scale * dz2;
  double hLocus = ellipsoidElevation(gp2[0], gp2[1], gp2[2], ell, desiredPrecision);
  i2g(hLocus, pos, achieved, st3);
  dir[0] = dx2;
  dir[1] = dy2;
  dir[2] = dz2;
  status = (st3 & ~USGSCSM_B200_ST_NONFINITE) | ((st1 | st2) & kErr);
  if (status & kErr) {
    status &= kErr;
    pos[0] = pos[1] = pos[2] = dir[0] = dir[1] = dir[2] = achieved = nan("");
  }
  if (!(finite3(pos[0], pos[1], pos[2]) && finite3(dir[0], dir[1], dir[2])))
    status |= USGSCSM_B200_ST_NONFINITE;
}

}  // namespace ub200
#endif
