// model_fr.cu -- host side of the framing-camera model: state -> packed device state.
//
// Mirrors UsgsAstroFrameSensorModel::populateModel (src/UsgsAstroFrameSensorModel.cpp:1362-1478)
// for the members the per-point math reads.  The model has one exposure, so the two
// camera-to-body matrices (calcRotationMatrix FR:2242-2286) are evaluated here, once.
#include <cmath>

#include "model.h"
#include "state_reader.h"

namespace ub200 {

void checkDistortionCoeffs(int type, size_t n);  // model_ls.cu

void repackFrame(Model &m) {
  FrameState &S = m.fr;
  for (int i = 0; i < 7; i++) S.params[i] = m.params[i] + 0.0;  // getValue(i, m_noAdjustments)
  for (int i = 0; i < 3; i++) S.rawParams[i] = m.params[i];
  {
    // calcRotationMatrix(m) FR:2242-2265 reads m_currentParameterValue directly
    double x = m.params[3], y = m.params[4], z = m.params[5], w = m.params[6];
    double norm = sqrt(x * x + y * y + z * z + w * w);
    x /= norm;
    y /= norm;
    w /= norm;
    z /= norm;
    double *r = S.mNorm;
    r[0] = w * w + x * x - y * y - z * z;
    r[1] = 2 * (x * y - w * z);
    r[2] = 2 * (w * y + x * z);
    r[3] = 2 * (x * y + w * z);
    r[4] = w * w - x * x + y * y - z * z;
    r[5] = 2 * (y * z - w * x);
    r[6] = 2 * (x * z - w * y);
    r[7] = 2 * (w * x + y * z);
    r[8] = w * w - x * x - y * y + z * z;
  }
  {
    // calcRotationMatrix(m, adjustments) FR:2267-2286: no normalisation
    double x = S.params[3], y = S.params[4], z = S.params[5], w = S.params[6];
    double *r = S.mRaw;
    r[0] = w * w + x * x - y * y - z * z;
    r[1] = 2 * (x * y - w * z);
    r[2] = 2 * (w * y + x * z);
    r[3] = 2 * (x * y + w * z);
    r[4] = w * w - x * x + y * y - z * z;
    r[5] = 2 * (y * z - w * x);
    r[6] = 2 * (x * z - w * y);
    r[7] = 2 * (w * x + y * z);
    r[8] = w * w - x * x - y * y + z * z;
  }
}

// groundToImage(groundPt, adj, ...) with adj[index] = delta (FR:844-873): getValue(i, adj) and
// the un-normalised matrix of the adjusted quaternion (calcRotationMatrix(m, adj) FR:2267-2286).
FrameState adjustedFrame(const Model &m, int index, double delta) {
  FrameState S = m.fr;
  for (int i = 0; i < 7; i++) S.params[i] = m.params[i] + (i == index ? delta : 0.0);
  double x = S.params[3], y = S.params[4], z = S.params[5], w = S.params[6];
  double *r = S.mRaw;
  r[0] = w * w + x * x - y * y - z * z;
  r[1] = 2 * (x * y - w * z);
  r[2] = 2 * (w * y + x * z);
  r[3] = 2 * (x * y + w * z);
  r[4] = w * w - x * x + y * y - z * z;
  r[5] = 2 * (y * z - w * x);
  r[6] = 2 * (x * z - w * y);
  r[7] = 2 * (w * x + y * z);
  r[8] = w * w - x * x - y * y + z * z;
  return S;
}

void loadFrame(Model &m, const StateReader &st) {
  FrameState &S = m.fr;
  S = FrameState{};
  FocalPlane &fp = S.fp;
  readEllipsoid(S.ell, st);
  fp.focalLength = st.getDouble("m_focalLength");
  fp.startingDetectorSample = st.getDouble("m_startingDetectorSample");
  fp.startingDetectorLine = st.getDouble("m_startingDetectorLine");
  fp.detectorSampleSumming = st.getInt("m_detectorSampleSumming");
  fp.detectorLineSumming = st.getInt("m_detectorLineSumming");
  fp.zDirection = 1.0;
  S.nLines = st.getInt("m_nLines");
  S.nSamples = st.getInt("m_nSamples");
  m.params = st.getVecDouble("m_currentParameterValue");
  if (m.params.size() < 7) throw StateError("m_currentParameterValue needs 7 entries");
  std::vector<double> ccdCenter = st.getVecDouble("m_ccdCenter");
  if (ccdCenter.size() < 2) throw StateError("m_ccdCenter needs 2 entries");
  fp.detectorLineOrigin = ccdCenter[0];    // computePixel(..., m_ccdCenter[1], m_ccdCenter[0], ...)
  fp.detectorSampleOrigin = ccdCenter[1];  // FR:186-189
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
  double determinant = fp.iTransS[2] * fp.iTransL[1] - fp.iTransS[1] * fp.iTransL[2];
  fp.p11 = fp.iTransS[2] / determinant;  // src/Utilities.cpp:143-147
  fp.p12 = -fp.iTransS[1] / determinant;
  fp.p21 = -fp.iTransL[2] / determinant;
  fp.p22 = fp.iTransL[1] / determinant;
  fp.lineSumming.b = fp.detectorLineSumming;
  fp.sampleSumming.b = fp.detectorSampleSumming;
  fillDivisors({&fp.lineSumming, &fp.sampleSumming});

  {
    std::vector<double> v = st.getVecDouble("m_spacecraftVelocity", {0.0, 0.0, 0.0});
    if (v.size() != 3) throw StateError("m_spacecraftVelocity needs 3 entries");
    for (int i = 0; i < 3; i++) m.fr.spacecraftVelocity[i] = v[i];
  }
  m.lineJitter = st.getVecDouble("m_lineJitter", {});
  m.sampleJitter = st.getVecDouble("m_sampleJitter", {});
  m.lineTimes = st.getVecDouble("m_lineTimes", {});
  // removeJitter throws INDEX_OUT_OF_RANGE for every point when the sizes differ (U:189-195)
  if (!m.lineTimes.empty() && m.lineJitter.size() != m.sampleJitter.size())
    throw StateError("Jitter coefficient vectors must be the same size.");
  S.numJitterCoeffs = (int)m.lineJitter.size();
  S.numLineTimes = (int)m.lineTimes.size();
  S.lineJitter = m.upload(m.lineJitter);
  S.sampleJitter = m.upload(m.sampleJitter);
  S.lineTimes = m.upload(m.lineTimes);
  repackFrame(m);
}

}  // namespace ub200
