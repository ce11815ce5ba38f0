// model_pf.cu -- host side of the push-frame model: state -> packed device state.
//
// Mirrors UsgsAstroPushFrameSensorModel::populateModel
// (src/UsgsAstroPushFrameSensorModel.cpp:141-298) and updateState (:605-676) for the members
// the per-point math reads, then builds the per-framelet exterior-orientation table the
// kernels read (sensor_state.h, LsState::frameletOrient).
#include <cmath>

#include "model.h"
#include "state_reader.h"

namespace ub200 {

namespace {

struct Scratch {
  double *p = nullptr;
  explicit Scratch(size_t doubles) { UB_CUDA(cudaMalloc(&p, doubles * sizeof(double))); }
  ~Scratch() { cudaFree(p); }
};

void buildFrameletOrientations(Model &m) {
  LsState &S = m.ls;
  S.frameletOrient = nullptr;  // the set-up kernel itself evaluates in-thread
  const size_t doubles = (size_t)(S.numFramelets + 2) * kFrameletOrientDoubles;
  if (!m.frameletOrient) {
    DeviceBuffer b;
    b.bytes = doubles * sizeof(double);
    UB_CUDA(cudaMalloc(&b.ptr, b.bytes));
    m.owned.push_back(b);
    m.frameletOrient = static_cast<double *>(b.ptr);
  }
  UB_CUDA(launchPfFrameletOrientations(S, m.frameletOrient, nullptr));
  UB_CUDA(cudaDeviceSynchronize());
  S.frameletOrient = m.frameletOrient;
}

// updateState PF:605-690.
void updateStatePf(Model &m) {
  LsState &S = m.ls;
  S.frameletOrient = nullptr;
  double line[2] = {S.nReducedLines / 2.0, S.nReducedLines / 2.0 + 1};
  double samp[2] = {S.nSamples / 2.0, S.nSamples / 2.0 + 1};
  double x[2], y[2], z[2];
  {
    Scratch d(2 * 5 + 1);
    double *dl = d.p, *ds = dl + 2, *dx = ds + 2, *dy = dx + 2, *dz = dy + 2, *dh = dz + 2;
    const double height = 0.0;
    UB_CUDA(cudaMemcpy(dl, line, sizeof(line), cudaMemcpyHostToDevice));
    UB_CUDA(cudaMemcpy(ds, samp, sizeof(samp), cudaMemcpyHostToDevice));
    UB_CUDA(cudaMemcpy(dh, &height, sizeof(double), cudaMemcpyHostToDevice));
    UB_CUDA(launchPfImageToGround(S, m.smCount, 2, dl, ds, nullptr, dh, 0, 0.001, dx, dy, dz,
                                  nullptr, nullptr, nullptr));
    UB_CUDA(cudaMemcpy(x, dx, sizeof(x), cudaMemcpyDeviceToHost));
    UB_CUDA(cudaMemcpy(y, dy, sizeof(y), cudaMemcpyDeviceToHost));
    UB_CUDA(cudaMemcpy(z, dz, sizeof(z), cudaMemcpyDeviceToHost));
  }
  m.referencePoint[0] = x[0];
  m.referencePoint[1] = y[0];
  m.referencePoint[2] = z[0];
  double dx = x[1] - x[0], dy = y[1] - y[0], dz = z[1] - z[0];
  S.gsd = sqrt((dx * dx + dy * dy + dz * dz) / 2.0);

  double sensor[3];
  {
    Scratch d(3);
    UB_CUDA(launchPfSensorPosition(S, 0.0, d.p, nullptr));  // getSensorPosition(0.0), PF:637
    UB_CUDA(cudaMemcpy(sensor, d.p, sizeof(sensor), cudaMemcpyDeviceToHost));
  }
  dx = sensor[0] - x[0];
  dy = sensor[1] - y[0];
  dz = sensor[2] - z[0];
  S.flyingHeight = sqrt(dx * dx + dy * dy + dz * dz);
  S.halfSwath = S.gsd * S.nSamples / 2.0;
  double fullImageTime = S.startingEphemerisTime + (S.numFramelets - 1) * S.interframeDelay +
                         S.exposureDuration;
  S.halfTime = fullImageTime / 2.0;
  fillDefaultCovariance(m.covariance, S.gsd);  // PF:660-690, same diagonal as the line-scan model
}

}  // namespace

void repackPushFrame(Model &m) {
  LsState &S = m.ls;
  S.hasAdj = 0;
  S.hasAttAdj = 0;
  for (int i = 0; i < 16; i++) {
    S.params[i] = m.params[i] + 0.0;  // getValue(i, _no_adjustment)
    if (S.params[i] != 0) S.hasAdj = 1;
    if (i >= 6 && i <= 14 && S.params[i] != 0) S.hasAttAdj = 1;
  }
  buildFrameletOrientations(m);
}

// The packed state groundToImage(ground_pt, adj, ...) sees with adj[index] = delta
// (PF:1141-1162).  The per-framelet orientation table depends on the parameters: the copy
// returned here has none (frameletOrient = nullptr, the kernels then evaluate in-thread);
// the caller may build one for it with launchPfFrameletOrientations.
LsState adjustedPushFrame(const Model &m, int index, double delta) {
  LsState S = m.ls;
  S.hasAdj = 0;
  S.hasAttAdj = 0;
  for (int i = 0; i < 16; i++) {
    S.params[i] = m.params[i] + (i == index ? delta : 0.0);
    if (S.params[i] != 0) S.hasAdj = 1;
    if (i >= 6 && i <= 14 && S.params[i] != 0) S.hasAttAdj = 1;
  }
  S.frameletOrient = nullptr;
  return S;
}

void loadPushFrame(Model &m, const StateReader &st) {
  LsState &S = m.ls;
  S = LsState{};
  m.frameletOrient = nullptr;
  S.nLines = st.getInt("m_nLines");
  S.nSamples = st.getInt("m_nSamples");
  S.exposureDuration = st.getDouble("m_exposureDuration");
  S.interframeDelay = st.getDouble("m_interframeDelay");
  (void)st.getInt("m_frameletHeight");
  S.frameletsFlipped = st.getBool("m_frameletsFlipped") ? 1 : 0;
  S.frameletsOrderReversed = st.getBool("m_frameletsOrderReversed") ? 1 : 0;
  S.startingEphemerisTime = st.getDouble("m_startingEphemerisTime");
  S.centerEphemerisTime = st.getDouble("m_centerEphemerisTime");

  readFocalPlane(S.fp, st, false);
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

  const int numLinesOverlap = st.getInt("m_numLinesOverlap");
  const int reducedFrameletHeight = st.getInt("m_reducedframeletHeight");  // sic, PF:279
  S.nReducedLines = st.getInt("m_nReducedLines");
  if (numLinesOverlap % 2 != 0)  // PF:289-295 throws INVALID_USE
    throw StateError("The number of overlapping lines among framelets must be even.");
  // int summedFrameletHeight = m_reducedFrameletHeight / m_detectorLineSumming (PF:721)
  const double q = reducedFrameletHeight / S.fp.detectorLineSumming;
  if (!(q >= 1.0 && q < 2147483648.0))
    throw StateError("m_reducedframeletHeight / m_detectorLineSumming must be at least 1");
  S.summedFrameletHeight = (int)q;
  S.numFramelets = S.nReducedLines / S.summedFrameletHeight;
  S.halfOverlap = numLinesOverlap / 2;

  // parameters first, without the table: updateState evaluates in-thread
  S.hasAdj = S.hasAttAdj = 0;
  for (int i = 0; i < 16; i++) {
    S.params[i] = m.params[i] + 0.0;
    if (S.params[i] != 0) S.hasAdj = 1;
    if (i >= 6 && i <= 14 && S.params[i] != 0) S.hasAttAdj = 1;
  }
  if (S.gsd == 1.0 && S.flyingHeight == 1000.0) updateStatePf(m);
  buildFrameletOrientations(m);
}

}  // namespace ub200
