// model_ls.cu -- host side of the line-scan model: state -> packed device state.
//
// Mirrors UsgsAstroLsSensorModel::populateModel (src/UsgsAstroLsSensorModel.cpp:141-293):
// copy the members the per-point math reads out of the state, then derive
// gsd / flyingHeight / halfSwath / halfTime (updateState, LS:649-735) and the projective
// seed (createProjectiveApproximation, LS:2745-2803).  The reference derives those with its
// own per-point calls; here the same points go through the batch kernels.
#include <cmath>
#include <cstring>

#include "model.h"
#include "state_reader.h"

namespace ub200 {

namespace {

struct DeviceScratch {
  double *p = nullptr;
  explicit DeviceScratch(size_t doubles) { UB_CUDA(cudaMalloc(&p, doubles * sizeof(double))); }
  ~DeviceScratch() { cudaFree(p); }
};

}  // namespace

void fillFocalPlaneInverse(FocalPlane &fp) {
  // computeDistortedFocalPlaneCoordinates src/Utilities.cpp:143-147
  double determinant = fp.iTransS[2] * fp.iTransL[1] - fp.iTransS[1] * fp.iTransL[2];
  fp.p11 = fp.iTransS[2] / determinant;
  fp.p12 = -fp.iTransS[1] / determinant;
  fp.p21 = -fp.iTransL[2] / determinant;
  fp.p22 = fp.iTransL[1] / determinant;
}

void fillDefaultCovariance(std::vector<double> &cov, double gsd);
namespace {

void fillEllipsoid(Ellipsoid &e, double a, double b) {
  e.majorAxis = a;
  e.minorAxis = b;
  e.eccSqr = 1.0 - b * b / a / a;  // src/Utilities.cpp:825
  e.ep2 = 1.0 - e.eccSqr;          // :826
}

// imageToGround of a handful of points through the batch kernel.
void imageToGroundFew(Model &m, int n, const double *line, const double *samp, double height,
                      double prec, double *x, double *y, double *z) {
  DeviceScratch d((size_t)n * 5 + 1);
  double *dl = d.p, *ds = dl + n, *dx = ds + n, *dy = dx + n, *dz = dy + n, *dh = dz + n;
  UB_CUDA(cudaMemcpy(dl, line, n * sizeof(double), cudaMemcpyHostToDevice));
  UB_CUDA(cudaMemcpy(ds, samp, n * sizeof(double), cudaMemcpyHostToDevice));
  UB_CUDA(cudaMemcpy(dh, &height, sizeof(double), cudaMemcpyHostToDevice));
  UB_CUDA(launchLsImageToGround(m.ls, m.smCount, n, dl, ds, nullptr, dh, 0, prec, dx, dy, dz,
                                nullptr, nullptr, nullptr));
  UB_CUDA(cudaMemcpy(x, dx, n * sizeof(double), cudaMemcpyDeviceToHost));
  UB_CUDA(cudaMemcpy(y, dy, n * sizeof(double), cudaMemcpyDeviceToHost));
  UB_CUDA(cudaMemcpy(z, dz, n * sizeof(double), cudaMemcpyDeviceToHost));
}

// updateState LS:649-735.
void updateStateLs(Model &m) {
  LsState &S = m.ls;
  const double lineCtr = S.nLines / 2.0;
  const double sampCtr = S.nSamples / 2.0;
  // default desiredPrecision of RasterGM::imageToGround is 0.001
  double line[2] = {lineCtr, lineCtr + 1}, samp[2] = {sampCtr, sampCtr + 1};
  double x[2], y[2], z[2];
  imageToGroundFew(m, 2, line, samp, 0.0, 0.001, x, y, z);
  m.referencePoint[0] = x[0];
  m.referencePoint[1] = y[0];
  m.referencePoint[2] = z[0];
  double dx = x[1] - x[0], dy = y[1] - y[0], dz = z[1] - z[0];
  S.gsd = sqrt((dx * dx + dy * dy + dz * dz) / 2.0);

  DeviceScratch d(4);
  double sensor[3];
  UB_CUDA(cudaMemcpy(d.p, &lineCtr, sizeof(double), cudaMemcpyHostToDevice));
  UB_CUDA(launchLsSensorPosition(S, 1, d.p, d.p + 1, nullptr));
  UB_CUDA(cudaMemcpy(sensor, d.p + 1, 3 * sizeof(double), cudaMemcpyDeviceToHost));
  dx = sensor[0] - x[0];
  dy = sensor[1] - y[0];
  dz = sensor[2] - z[0];
  S.flyingHeight = sqrt(dx * dx + dy * dy + dz * dz);
  S.halfSwath = S.gsd * S.nSamples / 2.0;
  double fullImageTime = m.intTimeStartTimes.back() - m.intTimeStartTimes.front() +
                         m.intTimes.back() * (S.nLines - m.intTimeLines.back());
  S.halfTime = fullImageTime / 2.0;
  fillDefaultCovariance(m.covariance, S.gsd);
}

}  // namespace

// Parameter covariance updateState hard-codes ("~1 pixel accuracy", LS:703-734, PF:660-690):
// a diagonal from the freshly derived gsd.  Read by the error-propagation entry points and
// written back by getModelState.
void fillDefaultCovariance(std::vector<double> &cov, double gsd) {
  const int n = 16;
  const double positionVariance = gsd * gsd;
  const double velocityVariance = positionVariance / 2.0;      // parameter time spans [0, 2]
  const double accelerationVariance = positionVariance / 4.0;
  cov.assign((size_t)n * n, 0.0);
  for (int i = 0; i < n; i++) {
    const bool rate = (i >= 3 && i <= 5) || (i >= 9 && i <= 11);
    const bool accel = i >= 12 && i <= 14;
    cov[(size_t)i * n + i] = accel ? accelerationVariance : (rate ? velocityVariance : positionVariance);
  }
}

namespace {

// createProjectiveApproximation LS:2745-2803
void createProjectiveApproximationLs(Model &m) {
  LsState &S = m.ls;
  S.useApproxInitTrans = 0;
  const int numPts = 9;
  const double u_factors[numPts] = {0.0, 0.0, 0.0, 0.5, 0.5, 0.5, 1.0, 1.0, 1.0};
  const double v_factors[numPts] = {0.0, 0.5, 1.0, 0.0, 0.5, 1.0, 0.0, 0.5, 1.0};

  double height;
  {
    DeviceScratch d(4);
    UB_CUDA(cudaMemcpy(d.p, m.referencePoint, 3 * sizeof(double), cudaMemcpyHostToDevice));
    UB_CUDA(launchEllipsoidElevation(S.ell, 1, d.p, 0.01, d.p + 3, nullptr));
    UB_CUDA(cudaMemcpy(&height, d.p + 3, sizeof(double), cudaMemcpyDeviceToHost));
  }
  if (std::isnan(height)) return;

  double line[2 * numPts], samp[2 * numPts], x[2 * numPts], y[2 * numPts], z[2 * numPts];
  for (int i = 0; i < numPts; i++) {
    line[i] = line[i + numPts] = u_factors[i] * (double)S.nLines;
    samp[i] = samp[i + numPts] = v_factors[i] * (double)S.nSamples;
  }
  imageToGroundFew(m, numPts, line, samp, height, 0.001, x, y, z);
  height += 100.0;
  imageToGroundFew(m, numPts, line + numPts, samp + numPts, height, 0.001, x + numPts,
                   y + numPts, z + numPts);
  // An error status (e.g. the LRO NAC zero denominator) is a throw in the reference, which
  // leaves m_useApproxInitTrans false (LS:287-292); such points come back NaN here.
  for (int i = 0; i < 2 * numPts; i++)
    if (std::isnan(x[i]) || std::isnan(y[i]) || std::isnan(z[i])) return;
  fitProjectiveTransform(2 * numPts, line, samp, x, y, z, S.projCoeffs);
  S.useApproxInitTrans = 1;
}

}  // namespace

void checkDistortionCoeffs(int type, size_t n) {
  // Size checks the reference makes per call (src/Distortion.cpp:120-127, 261-269,
  // 342-350, 408-416) are made once here, so that the kernels never see a short array.
  size_t need = 0;
  switch (type) {
    case DIST_RADIAL: need = 3; break;
    case DIST_TRANSVERSE: need = 20; break;  // 10 terms per axis at offsets 0 and size/2 (D:74-98)
    case DIST_KAGUYALISM: need = 10; break;
    case DIST_DAWNFC: need = 1; break;
    case DIST_LROLROCNAC: need = 1; break;
    case DIST_CAHVOR: need = 5; break;
    case DIST_RADTAN: need = 5; break;
    case DIST_KPLOSHADOWCAM: need = 1; break;
    default: need = 0; break;
  }
  if (n < need) throw StateError("m_opticalDistCoeffs is too short for the distortion model");
  if (n > (size_t)kMaxDistCoeffs) throw StateError("m_opticalDistCoeffs has more than 20 entries");
}


// lineScanTypes: the line-scan model reads the summing factors and m_zDirection as int
// (LS:170-190: a double in the state is truncated); push-frame reads all three as double
// (PF:166-191).
void readFocalPlane(FocalPlane &fp, const StateReader &st, bool lineScanTypes) {
  if (lineScanTypes) {
    fp.detectorSampleSumming = st.getInt("m_detectorSampleSumming");
    fp.detectorLineSumming = st.getInt("m_detectorLineSumming");
  } else {
    fp.detectorSampleSumming = st.getDouble("m_detectorSampleSumming");
    fp.detectorLineSumming = st.getDouble("m_detectorLineSumming");
  }
  fp.startingDetectorSample = st.getDouble("m_startingDetectorSample");
  fp.startingDetectorLine = st.getDouble("m_startingDetectorLine");
  fp.focalLength = st.getDouble("m_focalLength");
  fp.zDirection = lineScanTypes ? (double)st.getInt("m_zDirection") : st.getDouble("m_zDirection");
  fp.distortionType = st.getInt("m_distortionType");
  std::vector<double> coeffs = st.getVecDouble("m_opticalDistCoeffs");
  checkDistortionCoeffs(fp.distortionType, coeffs.size());
  fp.numDistCoeffs = (int)coeffs.size();
  for (int i = 0; i < kMaxDistCoeffs; i++) fp.distCoeffs[i] = i < fp.numDistCoeffs ? coeffs[i] : 0.0;
  std::vector<double> iTransS = st.getVecDouble("m_iTransS");
  std::vector<double> iTransL = st.getVecDouble("m_iTransL");
  if (iTransS.size() < 3 || iTransL.size() < 3) throw StateError("m_iTransS/m_iTransL need 3 entries");
  for (int i = 0; i < 3; i++) {
    fp.iTransS[i] = iTransS[i];
    fp.iTransL[i] = iTransL[i];
  }
  fp.detectorSampleOrigin = st.getDouble("m_detectorSampleOrigin");
  fp.detectorLineOrigin = st.getDouble("m_detectorLineOrigin");
  fillFocalPlaneInverse(fp);
  fp.lineSumming.b = fp.detectorLineSumming;
  fp.sampleSumming.b = fp.detectorSampleSumming;
  fillDivisors({&fp.lineSumming, &fp.sampleSumming});
}

void setEllipsoidRadii(Ellipsoid &e, double a, double b) { fillEllipsoid(e, a, b); }

void readEllipsoid(Ellipsoid &e, const StateReader &st) {
  fillEllipsoid(e, st.getDouble("m_majorAxis"), st.getDouble("m_minorAxis"));
}

// Shared by the line-scan and push-frame loaders: tables + exterior orientation grids.
void readOrientation(Model &m, const StateReader &st) {
  LsState &S = m.ls;
  m.platformFlag = st.getInt("m_platformFlag");
  const double dtEphem = st.getDouble("m_dtEphem");
  const double t0Ephem = st.getDouble("m_t0Ephem");
  const double dtQuat = st.getDouble("m_dtQuat");
  const double t0Quat = st.getDouble("m_t0Quat");
  const int numPositions = st.getInt("m_numPositions");  // counts doubles
  m.numQuaternionDoubles = st.getInt("m_numQuaternions");
  m.positions = st.getVecDouble("m_positions");
  m.velocities = st.getVecDouble("m_velocities");
  m.quaternions = st.getVecDouble("m_quaternions");
  if (numPositions < 0 || (size_t)numPositions > m.positions.size() ||
      (size_t)numPositions > m.velocities.size())
    throw StateError("m_numPositions exceeds the position/velocity tables");
  if (m.numQuaternionDoubles < 0 || (size_t)m.numQuaternionDoubles > m.quaternions.size())
    throw StateError("m_numQuaternions exceeds the quaternion table");
  // lagrangeInterp throws INDEX_OUT_OF_RANGE for fewer than 2 nodes (U:414-419): every
  // per-point call of such a model fails, so refuse it at load.
  if (numPositions / 3 < 2 || m.numQuaternionDoubles / 4 < 2)
    throw StateError("at least 2 ephemeris nodes and 2 quaternions are required");

  // LS:2494-2495, PF:1965-1966: order 8, or 4 when m_platformFlag == 0
  int nOrder = 8;
  if (m.platformFlag == 0) nOrder = 4;
  // LS:2142-2145, PF:1681-1684: the quaternion order also drops when m_numQuaternions < 6
  int nOrderQuat = 8;
  if (m.platformFlag == 0) nOrderQuat = 4;
  if (m.numQuaternionDoubles < 6 && nOrderQuat == 8) nOrderQuat = 4;

  S.positions = UniformTable{m.upload(m.positions), numPositions / 3, nOrder, t0Ephem, dtEphem};
  S.velocities = UniformTable{m.upload(m.velocities), numPositions / 3, nOrder, t0Ephem, dtEphem};
  S.quaternions =
      UniformTable{m.upload(m.quaternions), m.numQuaternionDoubles / 4, nOrderQuat, t0Quat, dtQuat};
  fillTableSteps({&S.positions, &S.velocities, &S.quaternions});
  std::vector<Divisor *> ds;
  appendLagrangeDivisors(S.lag, ds);
  fillDivisors(ds);
}

void repackLineScanner(Model &m) {
  LsState &S = m.ls;
  S.hasAdj = 0;
  S.hasAttAdj = 0;
  for (int i = 0; i < 16; i++) {
    // getValue(i, adj) = m_currentParameterValue[i] + adj[i], adj = 0 (LS:757-772)
    S.params[i] = m.params[i] + 0.0;
    if (i < 6 && S.params[i] != 0) S.hasAdj = 1;
    if (i >= 6 && i <= 14 && S.params[i] != 0) S.hasAttAdj = 1;
  }
}

// The packed state groundToImage(ground_pt, adj, ...) sees when adj[index] = delta and every
// other adjustment is 0: getValue(i, adj) = m_currentParameterValue[i] + adj[i] (LS:2129-2132).
LsState adjustedLineScanner(const Model &m, int index, double delta) {
  LsState S = m.ls;
  S.hasAdj = 0;
  S.hasAttAdj = 0;
  for (int i = 0; i < 16; i++) {
    S.params[i] = m.params[i] + (i == index ? delta : 0.0);
    if (i < 6 && S.params[i] != 0) S.hasAdj = 1;
    if (i >= 6 && i <= 14 && S.params[i] != 0) S.hasAttAdj = 1;
  }
  return S;
}

void loadLineScanner(Model &m, const StateReader &st) {
  LsState &S = m.ls;
  S = LsState{};
  S.nLines = st.getInt("m_nLines");
  S.nSamples = st.getInt("m_nSamples");
  m.intTimeLines = st.getVecDouble("m_intTimeLines");
  m.intTimeStartTimes = st.getVecDouble("m_intTimeStartTimes");
  m.intTimes = st.getVecDouble("m_intTimes");
  if (m.intTimeLines.empty() || m.intTimeStartTimes.size() < m.intTimeLines.size() ||
      m.intTimes.size() < m.intTimeLines.size())
    throw StateError("m_intTimeLines / m_intTimeStartTimes / m_intTimes are empty or ragged");
  S.numIntTimes = (int)m.intTimeLines.size();
  S.intLine0 = m.intTimeLines[0];
  S.intStart0 = m.intTimeStartTimes[0];
  S.intTime0 = m.intTimes[0];
  S.intTimeLines = m.upload(m.intTimeLines);
  S.intTimeStartTimes = m.upload(m.intTimeStartTimes);
  S.intTimes = m.upload(m.intTimes);
  S.startingEphemerisTime = st.getDouble("m_startingEphemerisTime");
  S.centerEphemerisTime = st.getDouble("m_centerEphemerisTime");

  readFocalPlane(S.fp, st, true);
  readEllipsoid(S.ell, st);
  readOrientation(m, st);

  std::vector<double> refPt = st.getVecDouble("m_referencePointXyz");
  if (refPt.size() < 3) throw StateError("m_referencePointXyz needs 3 entries");
  for (int i = 0; i < 3; i++) m.referencePoint[i] = refPt[i];
  S.gsd = st.getDouble("m_gsd");
  S.flyingHeight = st.getDouble("m_flyingHeight");
  S.halfSwath = st.getDouble("m_halfSwath");
  S.halfTime = st.getDouble("m_halfTime");
  m.params = st.getVecDouble("m_currentParameterValue");
  if (m.params.size() != 16) throw StateError("m_currentParameterValue needs 16 entries");
  repackLineScanner(m);

  if (S.gsd == 1.0 && S.flyingHeight == 1000.0) updateStateLs(m);
  createProjectiveApproximationLs(m);
}

}  // namespace ub200
