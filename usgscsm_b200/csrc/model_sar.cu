// model_sar.cu -- host side of the SAR model: state -> packed device state.
//
// Mirrors UsgsAstroSarSensorModel::populateModel (src/UsgsAstroSarSensorModel.cpp:232-326)
// for the members the per-point math reads.
#include <cmath>

#include "model.h"
#include "state_reader.h"

namespace ub200 {

void repackSar(Model &m) {
  for (int i = 0; i < 6; i++) m.sar.params[i] = m.params[i] + 0.0;  // getValue(i, m_noAdjustments)
}

// groundToImage(groundPt, adj, ...) with adj[index] = delta (SAR:1347-1371).
SarState adjustedSar(const Model &m, int index, double delta) {
  SarState S = m.sar;
  for (int i = 0; i < 6; i++) S.params[i] = m.params[i] + (i == index ? delta : 0.0);
  return S;
}

void loadSar(Model &m, const StateReader &st) {
  SarState &S = m.sar;
  S = SarState{};
  S.nLines = st.getInt("m_nLines");
  S.nSamples = st.getInt("m_nSamples");
  S.exposureDuration = st.getDouble("m_exposureDuration");
  S.scaledPixelWidth = st.getDouble("m_scaledPixelWidth");
  const std::string look = st.getString("m_lookDirection");
  if (look == "right")
    S.lookDirection = 1;
  else if (look == "left")
    S.lookDirection = 0;
  else
    throw StateError("Could not determine look direction from state");
  S.wavelength = st.getDouble("m_wavelength");
  S.startingEphemerisTime = st.getDouble("m_startingEphemerisTime");
  S.endingEphemerisTime = st.getDouble("m_endingEphemerisTime");
  readEllipsoid(S.ell, st);
  const double dtEphem = st.getDouble("m_dtEphem");
  const double t0Ephem = st.getDouble("m_t0Ephem");
  m.scaleCoeffs = st.getVecDouble("m_scaleConversionCoefficients");
  m.scaleTimes = st.getVecDouble("m_scaleConversionTimes");
  m.positions = st.getVecDouble("m_positions");
  m.velocities = st.getVecDouble("m_velocities");
  m.params = st.getVecDouble("m_currentParameterValue");
  if (m.params.size() < 6) throw StateError("m_currentParameterValue needs 6 entries");
  if (m.positions.size() < 3 || m.velocities.size() < 3)
    throw StateError("m_positions / m_velocities need at least one node");
  if (m.scaleCoeffs.size() < 4 || m.scaleTimes.empty())
    throw StateError("m_scaleConversionCoefficients / m_scaleConversionTimes are empty");

  S.positions = UniformTable{m.upload(m.positions), (int)(m.positions.size() / 3), 8, t0Ephem, dtEphem};
  S.velocities = UniformTable{m.upload(m.velocities), (int)(m.velocities.size() / 3), 8, t0Ephem, dtEphem};
  // getRangeCoefficients SAR:1941-1946: the step divides by the number of sets, not by N - 1
  const int numSets = (int)(m.scaleCoeffs.size() / 4);
  const double dtScale = (m.scaleTimes.back() - m.scaleTimes[0]) / (m.scaleCoeffs.size() / 4);
  S.scale = UniformTable{m.upload(m.scaleCoeffs), numSets, 8, m.scaleTimes[0], dtScale};
  fillTableSteps({&S.positions, &S.velocities, &S.scale});
  std::vector<Divisor *> ds;
  appendLagrangeDivisors(S.lag, ds);
  S.exposure.b = S.exposureDuration;
  S.pixelWidth.b = S.scaledPixelWidth;
  ds.push_back(&S.exposure);
  ds.push_back(&S.pixelWidth);
  fillDivisors(ds);
  repackSar(m);
}

}  // namespace ub200
