// sensor_state.h -- packed, device-ready model states (plain PODs shared by host and kernels).
//
// A packed state is the subset of a sensor model's members that the per-point math
// reads (SURVEY.md §8b "Packed device state"): scalars by value, tables as device
// pointers.  The struct is passed to the kernels as a __grid_constant__ parameter, so its
// scalars live in the constant bank and FP64 instructions read them as immediate
// operands; the tables live in HBM/L2 and are staged into shared memory by the kernels
// when they fit.
//
// Reference members mirrored here:
//   line-scan  include/usgscsm/UsgsAstroLsSensorModel.h:77-131,983-989
//   push-frame include/usgscsm/UsgsAstroPushFrameSensorModel.h (framelet fields)
//   frame      include/usgscsm/UsgsAstroFrameSensorModel.h
//   SAR        include/usgscsm/UsgsAstroSarSensorModel.h
#ifndef USGSCSM_B200_SENSOR_STATE_H_
#define USGSCSM_B200_SENSOR_STATE_H_

#include <stdint.h>

namespace ub200 {

// include/usgscsm/Distortion.h:10-11 (same numeric values: they are what m_distortionType
// holds in a serialised model state).
enum DistortionKind : int {
  DIST_RADIAL = 0,
  DIST_TRANSVERSE = 1,
  DIST_KAGUYALISM = 2,
  DIST_DAWNFC = 3,
  DIST_LROLROCNAC = 4,
  DIST_CAHVOR = 5,
  DIST_LUNARORBITER = 6,
  DIST_RADTAN = 7,
  DIST_KPLOSHADOWCAM = 8,
};

constexpr int kMaxDistCoeffs = 20;

// A divisor that is constant for a launch, with the reciprocal the compiler's own IEEE
// division would derive from it (device_math.cuh, refinedReciprocal), evaluated once per
// model by a set-up kernel.  divideBy(a, d) == a / d.b bit for bit, at 3 FP64 operations.
struct Divisor {
  double b;
  double r;
};

// The factorial constants of the Lagrange weights (src/Utilities.cpp:455-505).
struct LagrangeDivisors {
  Divisor d5040, d720, d240, d144;  // order 8
  Divisor d120, d24, d12;           // order 6
  Divisor d6;                       // order 4 (the other order-4 divisor, 2.0, is a * 0.5)
};

// Focal-plane <-> detector affine, shared by all optical models.
struct FocalPlane {
  double iTransS[3];
  double iTransL[3];
  double detectorSampleOrigin;
  double detectorLineOrigin;
  double detectorSampleSumming;
  double detectorLineSumming;
  double startingDetectorSample;
  double startingDetectorLine;
  double focalLength;
  double zDirection;  // LS/PF only
  int distortionType;
  int numDistCoeffs;
  double distCoeffs[kMaxDistCoeffs];
  // Inverse of the 2x2 part of (iTransL, iTransS), evaluated once on the host with the
  // exact expressions of computeDistortedFocalPlaneCoordinates (src/Utilities.cpp:143-147;
  // IEEE division is deterministic, so this equals the per-call value bit for bit).
  double p11, p12, p21, p22;
  Divisor lineSumming;    // detectorLineSumming
  Divisor sampleSumming;  // detectorSampleSumming
};

// Ellipsoid constants of computeEllipsoidElevation (src/Utilities.cpp:825-826), evaluated
// once on the host with the same expressions.
struct Ellipsoid {
  double majorAxis;
  double minorAxis;
  double eccSqr;  // 1.0 - b * b / a / a
  double ep2;     // 1.0 - eccSqr
};

// Uniformly sampled table of `vlen`-vectors for lagrangeInterp (src/Utilities.cpp:409-513).
struct UniformTable {
  const double *values;  // device pointer, numNodes * vlen doubles
  int numNodes;
  int maxOrder;  // i_order passed by the model (8, or 4: LS:2142-2145, LS:2494-2495)
  double t0;
  double dt;
  double rdt;  // refinedReciprocal(dt)
};

// Line-scan and push-frame models share the time-dependent exterior orientation.
struct LsState {
  FocalPlane fp;
  LagrangeDivisors lag;
  int nLines;
  int nSamples;
  int numIntTimes;        // m_intTimeLines.size()
  int useApproxInitTrans; // LS:2688-2736
  // Line-scan: any of parameters 0..5 (position / velocity bias and rate) != 0.  The reference
  // takes its adjusted-position path when ANY parameter is non-zero (LS:2505-2509), but with
  // parameters 0..5 all zero that path adds products with exact zeros to the nominal position
  // (LS:2567-2583): same bits, so the frame and the velocity interpolation are only needed for
  // these six.  Push-frame: any parameter != 0 (its position routine has no shortcut).
  int hasAdj;
  int hasAttAdj;          // any of parameters 6..14 != 0 (Euler angles not identically 0)
  Ellipsoid ell;
  double gsd;
  double flyingHeight;
  double halfSwath;
  double halfTime;
  double projCoeffs[14];
  double params[16];
  // first line-rate segment inline (every fixture has exactly one)
  double intLine0, intStart0, intTime0;
  const double *intTimeLines;      // device, numIntTimes
  const double *intTimeStartTimes; // device
  const double *intTimes;          // device
  UniformTable positions;          // vlen 3
  UniformTable velocities;         // vlen 3 (same grid as positions)
  UniformTable quaternions;        // vlen 4
  // push-frame only (PF:715-778, PF:1258-1276, PF:1726-1824)
  int nReducedLines;          // m_nReducedLines (the push-frame image height)
  int summedFrameletHeight;   // int(m_reducedFrameletHeight / m_detectorLineSumming), PF:721
  int numFramelets;           // m_nReducedLines / summedFrameletHeight (integer division)
  int halfOverlap;            // m_numLinesOverlap / 2 (integer division), PF:764
  int frameletsFlipped;
  int frameletsOrderReversed;
  double startingEphemerisTime;
  double centerEphemerisTime;
  double exposureDuration;
  double interframeDelay;
  // Exterior orientation per framelet time.  A push-frame image has one exposure time per
  // framelet (PF:1258-1276), so the adjusted sensor position, the camera-to-body matrix
  // and the attitude correction take only numFramelets + 2 distinct values (time framelet
  // numbers -1 .. numFramelets).  A set-up kernel evaluates them once with the per-point
  // device functions; entry k + 1 holds pos[3], camToBody[9], attCorr[9] (kFrameletOrientDoubles).
  const double *frameletOrient;  // device, (numFramelets + 2) * kFrameletOrientDoubles; may be null
};

constexpr int kFrameletOrientDoubles = 21;

// Frame model (FR:137-372, FR:2242-2385).
struct FrameState {
  FocalPlane fp;  // detector origins = m_ccdCenter[1] (sample), m_ccdCenter[0] (line); zDirection unused
  int nLines;
  int nSamples;
  int numJitterCoeffs;  // m_lineJitter.size() (== m_sampleJitter.size())
  int numLineTimes;     // m_lineTimes.size(); 0 = no rolling-shutter jitter correction
  Ellipsoid ell;
  double params[7];  // x, y, z, qx, qy, qz, qw (m_currentParameterValue + 0.0)
  double rawParams[3];  // m_currentParameterValue[0..2] as stored (imageToRemoteImagingLocus FR:540)
  double spacecraftVelocity[3];  // m_spacecraftVelocity (getSensorVelocity FR:749-795)
  // Camera-to-body matrices, evaluated once on the host with the reference's expressions
  // (IEEE arithmetic without contraction on both sides, hence the same bits):
  double mNorm[9];  // calcRotationMatrix(m) FR:2242-2265: quaternion normalised (imageToGround)
  double mRaw[9];   // calcRotationMatrix(m, adj) FR:2267-2286: NOT normalised (groundToImage)
  const double *lineJitter;    // device
  const double *sampleJitter;  // device
  const double *lineTimes;     // device
};

// SAR model (SAR:541-861, SAR:1196-1300, SAR:1936-1972).
struct SarState {
  int nLines;
  int nSamples;
  int lookDirection;  // 0 = LEFT, 1 = RIGHT (include/usgscsm/UsgsAstroSarSensorModel.h)
  int pad0;
  Ellipsoid ell;
  double exposureDuration;
  double scaledPixelWidth;
  double wavelength;
  double startingEphemerisTime;
  double endingEphemerisTime;
  double params[6];  // position and velocity offsets (m_currentParameterValue + 0.0)
  // positions.numNodes = m_positions.size() / 3, velocities.numNodes = m_velocities.size() / 3
  // (SAR:1199, 1269): a table with a single node is used as a constant.
  UniformTable positions;   // vlen 3, order 8
  UniformTable velocities;  // vlen 3, order 8, same t0 / dt
  // m_scaleConversionCoefficients, 4 per node; t0 = m_scaleConversionTimes[0],
  // dt = (m_scaleConversionTimes.back() - t0) / numNodes  (sic: /N, SAR:1944-1945)
  UniformTable scale;       // vlen 4, order 8
  LagrangeDivisors lag;
  Divisor exposure;    // exposureDuration
  Divisor pixelWidth;  // scaledPixelWidth
};

}  // namespace ub200
#endif
