// fr_kernels.cu -- framing-camera groundToImage / imageToGround, one thread per point, FP64.
//
// Replaces UsgsAstroFrameSensorModel::groundToImage (src/UsgsAstroFrameSensorModel.cpp:137-232)
// and ::imageToGround (:280-372) with losEllipsoidIntersect (:2288-2345) and the
// rolling-shutter jitter polynomial removeJitter / addJitter (src/Utilities.cpp:182-282).
//
// A frame has one exposure: position and attitude are model constants, so the two rotation
// matrices are evaluated once on the host (sensor_state.h) and the per-point work is ~100
// flop against 44-52 B of HBM traffic: this is the one HBM-bound kernel pair of the engine
// (SURVEY.md §8d).  I/O is SoA, coalesced, streaming.
#include <cstdint>
#include <cstdlib>
#include <initializer_list>

#include "kernels.h"
#include "device_math.cuh"

namespace ub200 {

namespace {

constexpr int kThreads = 256;
// resident blocks of 256 threads per SM the register allocation is held to (measured best:
// profiles/r2_ab_c3_occupancy.txt)
#ifndef UB_FR_G2I_MINB
#define UB_FR_G2I_MINB 4
#endif
#ifndef UB_FR_I2G_MINB
#define UB_FR_I2G_MINB 5
#endif

// removeJitter U:182-213.  Coefficients are in descending power, no constant term.
UB_DEV void removeJitter(const FrameState &S, double line, double sample, double &dejitteredLine,
                         double &dejitteredSample) {
  double lineJitter = 0;
  double sampleJitter = 0;
  int idx = max(min(truncToInt(round(line)), S.numLineTimes), 1) - 1;
  double time = __ldg(S.lineTimes + idx);
  for (int n = 0; n < S.numJitterCoeffs; n++) {
    double timeTerm = pow(time, (double)(S.numJitterCoeffs - n));
    lineJitter += __ldg(S.lineJitter + n) * timeTerm;
    sampleJitter += __ldg(S.sampleJitter + n) * timeTerm;
  }
  dejitteredLine = line - lineJitter;
  dejitteredSample = sample - sampleJitter;
}

// addJitter U:244-282
UB_DEV void addJitter(const FrameState &S, double line, double sample, double tolerance,
                      int maxIts, double &jitteredLine, double &jitteredSample) {
  int iteration = 0;
  double dejitteredLine = line - 1;
  double dejitteredSample = sample - 1;
  double currentLine = line;
  double currentSample = sample;
  while (iteration < maxIts) {
    removeJitter(S, currentLine, currentSample, dejitteredLine, dejitteredSample);
    if (fabs(dejitteredLine - line) < tolerance && fabs(dejitteredSample - sample) < tolerance)
      break;
    currentLine = line + currentLine - dejitteredLine;
    currentSample = sample + currentSample - dejitteredSample;
    iteration++;
  }
  jitteredLine = currentLine;
  jitteredSample = currentSample;
}

// removeJitter / addJitter over arrays (known-answer replay, usgscsm_b200_jitter)
__global__ void jitterKernel(const __grid_constant__ FrameState S, int remove, double tolerance,
                             int maxIts, long n, const double *__restrict__ line,
                             const double *__restrict__ sample, double *__restrict__ outLine,
                             double *__restrict__ outSample) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double l, s;
    if (remove)
      removeJitter(S, line[i], sample[i], l, s);
    else
      addJitter(S, line[i], sample[i], tolerance, maxIts, l, s);
    outLine[i] = l;
    outSample[i] = s;
  }
}

// DIST >= 0 / JITTER = false: the launcher knows the model's distortion type and that it has no
// jitter tables, and only that arithmetic is compiled in (the same operations on the same values).
template <int DIST = -1, bool JITTER = true>
UB_DEV void frGroundToImagePoint(const FrameState &S, double x, double y, double z,
                                 double desiredPrecision, double &outLine, double &outSamp,
                                 uint32_t &status) {
  const double *m = S.mRaw;  // m[r][c] = m[3 * r + c]
  double xo = x - S.params[0];
  double yo = y - S.params[1];
  double zo = z - S.params[2];
  double f = S.fp.focalLength;
  double denom = m[2] * xo + m[5] * yo + m[8] * zo;
  const double rdenom = refinedReciprocal(denom);  // one reciprocal for the two quotients
  double undistortedx = divideBy(f * (m[0] * xo + m[3] * yo + m[6] * zo), denom, rdenom);
  double undistortedy = divideBy(f * (m[1] * xo + m[4] * yo + m[7] * zo), denom, rdenom);
  double distortedX, distortedY;
  applyDistortion<DIST>(undistortedx, undistortedy, distortedX, distortedY, S.fp, 1.0E-6, status);
  double line, sample;
  computePixel(distortedX, distortedY, S.fp, line, sample);
  if (JITTER && S.numLineTimes) {
    double jitteredLine, jitteredSample;
    addJitter(S, line, sample, desiredPrecision, 20, jitteredLine, jitteredSample);
    line = jitteredLine;
    sample = jitteredSample;
  }
  outLine = line;
  outSamp = sample;
}

template <int DIST = -1, bool JITTER = true>
UB_DEV void frImageToGroundPoint(const FrameState &S, double line, double sample, double height,
                                 double &x, double &y, double &z, uint32_t &status) {
  const double *m = S.mNorm;
  if (JITTER && S.numLineTimes) {
    double dejitteredLine, dejitteredSample;
    removeJitter(S, line, sample, dejitteredLine, dejitteredSample);
    line = dejitteredLine;
    sample = dejitteredSample;
  }
  double x_camera, y_camera;
  distortedFocalPlaneCoordinates(line, sample, S.fp, x_camera, y_camera);
  double undistortedX, undistortedY;
  removeDistortion<DIST>(x_camera, y_camera, undistortedX, undistortedY, S.fp, 1.0E-6, status);
  const double f = S.fp.focalLength;
  double xl = m[0] * undistortedX + m[1] * undistortedY - m[2] * -f;
  double yl = m[3] * undistortedX + m[4] * undistortedY - m[5] * -f;
  double zl = m[6] * undistortedX + m[7] * undistortedY - m[8] * -f;
  double xc = S.params[0], yc = S.params[1], zc = S.params[2];

  // losEllipsoidIntersect FR:2288-2345: always the (-bt - sqrt) root
  double ap = S.ell.majorAxis + height;
  double bp = S.ell.minorAxis + height;
  // sphere: ap and bp are the same double, and x / x == 1 for any finite non-zero x
  double ap2 = ap * ap;
  double k = (S.ell.majorAxis == S.ell.minorAxis && ap2 > 0.0 && ap2 < CUDART_INF) ? 1.0 : ap2 / (bp * bp);
  double at = xl * xl + yl * yl + k * zl * zl;
  double bt = 2.0 * (xl * xc + yl * yc + k * zl * zc);
  double ct = xc * xc + yc * yc + k * zc * zc - ap * ap;
  double quadTerm = bt * bt - 4.0 * at * ct;
  if (0.0 > quadTerm) {
    quadTerm = 0.0;
    status |= USGSCSM_B200_ST_NO_INTERSECTION;
  }
  double scale = (-bt - sqrt(quadTerm)) / (2.0 * at);
  x = xc + scale * xl;
  y = yc + scale * yl;
  z = zc + scale * zl;
  if (status & USGSCSM_B200_ST_ERROR_ALGORITHM) {  // removeDistortion threw (D:357-365)
    status = USGSCSM_B200_ST_ERROR_ALGORITHM;
    x = y = z = nan("");
  }
  if (!finite3(x, y, z)) status |= USGSCSM_B200_ST_NONFINITE;
}

template <int DIST, bool JITTER>
UB_DEV void frGroundToImageChecked(const FrameState &S, double gx, double gy, double gz, double prec,
                                   double &l, double &s, uint32_t &st) {
  st = 0;
  frGroundToImagePoint<DIST, JITTER>(S, gx, gy, gz, prec, l, s, st);
  if (!(isfinite(l) && isfinite(s))) st |= USGSCSM_B200_ST_NONFINITE;
  // getValidImageRange FR:600-608: (startingDetectorLine, startingDetectorSample) .. (nLines, nSamples)
  if (!(l >= S.fp.startingDetectorLine && l <= (double)S.nLines &&
        s >= S.fp.startingDetectorSample && s <= (double)S.nSamples))
    st |= USGSCSM_B200_ST_OUT_OF_IMAGE;
}

template <int DIST, bool JITTER, int VEC>
__global__ void __launch_bounds__(kThreads, UB_FR_G2I_MINB)
    frGroundToImageKernel(const __grid_constant__ FrameState S, long n, const double *__restrict__ x,
                          const double *__restrict__ y, const double *__restrict__ z, double prec,
                          double *__restrict__ line, double *__restrict__ samp,
                          double *__restrict__ achieved, uint32_t *__restrict__ status) {
  typedef PointPack<VEC> P;
  typedef typename P::D D;
  typedef typename P::U U;
  const long stride = (long)gridDim.x * blockDim.x;
  const long first = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long packs = n / VEC;
  const double never[2] = {nan(""), nan("")};  // the frame model never writes achievedPrecision
  for (long p = first; p < packs; p += stride) {
    const D gx = __ldcs(reinterpret_cast<const D *>(x) + p);
    const D gy = __ldcs(reinterpret_cast<const D *>(y) + p);
    const D gz = __ldcs(reinterpret_cast<const D *>(z) + p);
    double l[VEC], s[VEC];
    uint32_t st[VEC];
#pragma unroll
    for (int k = 0; k < VEC; k++)
      frGroundToImageChecked<DIST, JITTER>(S, P::get(gx, k), P::get(gy, k), P::get(gz, k), prec, l[k],
                                           s[k], st[k]);
    __stcs(reinterpret_cast<D *>(line) + p, P::make(l));
    __stcs(reinterpret_cast<D *>(samp) + p, P::make(s));
    if (achieved) __stcs(reinterpret_cast<D *>(achieved) + p, P::make(never));
    if (status) __stcs(reinterpret_cast<U *>(status) + p, P::makeStatus(st));
  }
  if (VEC > 1 && first == 0)
    for (long i = packs * VEC; i < n; i++) {
      double l, s;
      uint32_t st;
      frGroundToImageChecked<DIST, JITTER>(S, x[i], y[i], z[i], prec, l, s, st);
      line[i] = l;
      samp[i] = s;
      if (achieved) achieved[i] = never[0];
      if (status) status[i] = st;
    }
}

template <bool GRID, int DIST, bool JITTER, int VEC>
__global__ void __launch_bounds__(kThreads, UB_FR_I2G_MINB)
    frImageToGroundKernel(const __grid_constant__ FrameState S, long n, const double *__restrict__ line,
                          const double *__restrict__ samp, GridSpec g,
                          const double *__restrict__ height, int heightStride,
                          double *__restrict__ x, double *__restrict__ y, double *__restrict__ z,
                          double *__restrict__ achieved, uint32_t *__restrict__ status) {
  typedef PointPack<VEC> P;
  typedef typename P::D D;
  typedef typename P::U U;
  const long stride = (long)gridDim.x * blockDim.x;
  const long first = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long packs = n / VEC;
  const double never[2] = {nan(""), nan("")};
  const double oneHeight = heightStride ? 0.0 : __ldg(height);
  // dense grid (its rows hold a whole number of packs, the launcher checks): (row, column) of
  // the thread's pack, advanced by the stride's (rows, columns) instead of a division per point
  long r = 0, strideRows = 0;
  long c = 0, strideCols = 0;
  if (GRID) {
    const long perRow = g.nSamps / VEC;
    r = first / perRow;
    c = (first - r * perRow) * VEC;
    strideRows = stride / perRow;
    strideCols = (stride - strideRows * perRow) * VEC;
  }
  for (long p = first; p < packs; p += stride) {
    D lv, sv, hv;
    if (!GRID) {
      lv = __ldcs(reinterpret_cast<const D *>(line) + p);
      sv = __ldcs(reinterpret_cast<const D *>(samp) + p);
    }
    if (heightStride) hv = __ldcs(reinterpret_cast<const D *>(height) + p);
    double px[VEC], py[VEC], pz[VEC];
    uint32_t st[VEC];
#pragma unroll
    for (int k = 0; k < VEC; k++) {
      double l, sm;
      if (GRID) {
        l = g.line0 + (double)(g.row0 + r) * g.dline;
        sm = g.samp0 + (double)(c + k) * g.dsamp;
      } else {
        l = P::get(lv, k);
        sm = P::get(sv, k);
      }
      st[k] = 0;
      frImageToGroundPoint<DIST, JITTER>(S, l, sm, heightStride ? P::get(hv, k) : oneHeight, px[k],
                                         py[k], pz[k], st[k]);
    }
    if (GRID) {
      r += strideRows;
      c += strideCols;
      if (c >= g.nSamps) {
        c -= g.nSamps;
        r++;
      }
    }
    __stcs(reinterpret_cast<D *>(x) + p, P::make(px));
    __stcs(reinterpret_cast<D *>(y) + p, P::make(py));
    __stcs(reinterpret_cast<D *>(z) + p, P::make(pz));
    if (achieved) __stcs(reinterpret_cast<D *>(achieved) + p, P::make(never));
    if (status) __stcs(reinterpret_cast<U *>(status) + p, P::makeStatus(st));
  }
  if (VEC > 1 && !GRID && first == 0)
    for (long i = packs * VEC; i < n; i++) {
      double px, py, pz;
      uint32_t st = 0;
      frImageToGroundPoint<DIST, JITTER>(S, line[i], samp[i], heightStride ? height[i] : oneHeight, px,
                                         py, pz, st);
      x[i] = px;
      y[i] = py;
      z[i] = pz;
      if (achieved) achieved[i] = never[0];
      if (status) status[i] = st;
    }
}

// computeGroundPartials FR:955-1003: the step is the ground size of a pixel, from
// imageToGround (height 0) of the pixel one line and one sample further (one back if that
// collapses); then forward differences of groundToImage at default precision.
__global__ void __launch_bounds__(kThreads)
    frGroundPartialsKernel(const __grid_constant__ FrameState S, long n, const double *__restrict__ x,
                           const double *__restrict__ y, const double *__restrict__ z,
                           double *__restrict__ partials, uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double g[3] = {__ldcs(x + i), __ldcs(y + i), __ldcs(z + i)};
    double l[4], s[4];
    uint32_t st = 0;
    frGroundToImagePoint(S, g[0], g[1], g[2], 0.001, l[0], s[0], st);
    double nx, ny, nz;
    frImageToGroundPoint(S, l[0] + 1, s[0] + 1, 0.0, nx, ny, nz, st);
    double dx = nx - g[0], dy = ny - g[1], dz = nz - g[2];
    double delta = sqrt((dx * dx + dy * dy + dz * dz) / 2.0);
    if (delta < 1e-10) {
      frImageToGroundPoint(S, l[0] - 1, s[0] - 1, 0.0, nx, ny, nz, st);
      dx = nx - g[0];
      dy = ny - g[1];
      dz = nz - g[2];
      delta = sqrt((dx * dx + dy * dy + dz * dz) / 2.0);
    }
#pragma unroll 1
    for (int k = 1; k < 4; k++) {
      double q[3] = {g[0], g[1], g[2]};
      q[k - 1] = g[k - 1] + delta;
      frGroundToImagePoint(S, q[0], q[1], q[2], 0.001, l[k], s[k], st);
    }
    double p[6];
    p[0] = (l[1] - l[0]) / delta;
    p[3] = (s[1] - s[0]) / delta;
    p[1] = (l[2] - l[0]) / delta;
    p[4] = (s[2] - s[0]) / delta;
    p[2] = (l[3] - l[0]) / delta;
    p[5] = (s[3] - s[0]) / delta;
    storePartials(partials, status, i, p, st);
  }
}

// imageToRemoteImagingLocus FR:488-543: no de-jittering; point = the RAW parameters 0..2
// (m_currentParameterValue), direction = unit(m * (ux, uy, f)); achievedPrecision = 0.
UB_DEV void frRemoteLocusPoint(const FrameState &S, double line, double samp, double pos[3],
                               double dir[3], uint32_t &st) {
  const double *m = S.mNorm;
  double focalPlaneX, focalPlaneY;
  distortedFocalPlaneCoordinates(line, samp, S.fp, focalPlaneX, focalPlaneY);
  double ux, uy;
  removeDistortion(focalPlaneX, focalPlaneY, ux, uy, S.fp, 1.0E-6, st);
  const double f = S.fp.focalLength;
  double lookB[3] = {m[0] * ux + m[1] * uy + m[2] * f, m[3] * ux + m[4] * uy + m[5] * f,
                     m[6] * ux + m[7] * uy + m[8] * f};
  double mag = sqrt(lookB[0] * lookB[0] + lookB[1] * lookB[1] + lookB[2] * lookB[2]);
  dir[0] = lookB[0] / mag;
  dir[1] = lookB[1] / mag;
  dir[2] = lookB[2] / mag;
  pos[0] = S.rawParams[0];
  pos[1] = S.rawParams[1];
  pos[2] = S.rawParams[2];
}

UB_DEV void frFinishLocus(double pos[3], double dir[3], uint32_t &st) {
  if (st & USGSCSM_B200_ST_ERROR_ALGORITHM) {
    st = USGSCSM_B200_ST_ERROR_ALGORITHM;
    pos[0] = pos[1] = pos[2] = dir[0] = dir[1] = dir[2] = nan("");
  }
  if (!(finite3(pos[0], pos[1], pos[2]) && finite3(dir[0], dir[1], dir[2])))
    st |= USGSCSM_B200_ST_NONFINITE;
}

__global__ void __launch_bounds__(kThreads)
    frRemoteLocusKernel(const __grid_constant__ FrameState S, long n, const double *__restrict__ line,
                        const double *__restrict__ samp, LocusOut o) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t st = 0;
    double pos[3], dir[3];
    frRemoteLocusPoint(S, __ldcs(line + i), __ldcs(samp + i), pos, dir, st);
    frFinishLocus(pos, dir, st);
    storeLocus(o, i, pos, dir, 0.0, st);
  }
}

// imageToProximateImagingLocus FR:424-462: the point of the remote locus' line closest to the
// ground point: point + dot(direction, ground - point) * direction.
__global__ void __launch_bounds__(kThreads)
    frProximateLocusKernel(const __grid_constant__ FrameState S, long n, const double *__restrict__ line,
                           const double *__restrict__ samp, const double *__restrict__ x,
                           const double *__restrict__ y, const double *__restrict__ z, LocusOut o) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t st = 0;
    double pos[3], dir[3];
    frRemoteLocusPoint(S, __ldcs(line + i), __ldcs(samp + i), pos, dir, st);
    double toGround[3] = {__ldcs(x + i) - pos[0], __ldcs(y + i) - pos[1], __ldcs(z + i) - pos[2]};
    double d = dir[0] * toGround[0] + dir[1] * toGround[1] + dir[2] * toGround[2];
    double pt[3] = {pos[0] + d * dir[0], pos[1] + d * dir[1], pos[2] + d * dir[2]};
    frFinishLocus(pt, dir, st);
    storeLocus(o, i, pt, dir, 0.0, st);
  }
}

// A few waves of as many blocks as are resident at once for THIS kernel (its register count
// decides: 3 x 256 threads per SM at 80 registers, 4 at 64); the grid-stride loop splits the
// points evenly over them.
template <class K>
int frBlocks(K kernel, long n, int smCount) {
  static const int perSm = [&] {
    int b = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, kernel, kThreads, 0) != cudaSuccess || b < 1)
      b = 4;
    return b;
  }();
  long blocksNeeded = (n + kThreads - 1) / kThreads;
  // USGSCSM_B200_FR_WAVES: blocks per resident slot.  Four short waves balance the tail of a
  // launch that lasts 0.15 ms better than one (single 4096^2 image: +6 %; 16 and up cost again)
  static const int waves = [] {
    const char *v = getenv("USGSCSM_B200_FR_WAVES");
    const int w = v && *v ? atoi(v) : 4;
    return w < 1 ? 1 : w;
  }();
  long persistent = (long)smCount * perSm * waves;
  long b = blocksNeeded < persistent ? blocksNeeded : persistent;
  return (int)(b < 1 ? 1 : b);
}

// getImageTime FR:665-673 (0.0: the whole image is acquired at once), getSensorPosition
// (ImageCoord) FR:689-700 (m_currentParameterValue[0..2]), getSensorVelocity(ImageCoord)
// FR:749-765 (m_spacecraftVelocity; csm::Error::BOUNDS outside the image).
__global__ void frSensorStateKernel(const __grid_constant__ FrameState S, long n,
                                    const double *__restrict__ line,
                                    const double *__restrict__ samp, SensorStateOut o) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double pos[3] = {S.rawParams[0], S.rawParams[1], S.rawParams[2]};
    double vel[3] = {S.spacecraftVelocity[0], S.spacecraftVelocity[1], S.spacecraftVelocity[2]};
    uint32_t st = 0;
    if (o.atTime) {
      // getSensorPosition(double) / getSensorVelocity(double) FR:707-745, 769-795: the only valid
      // image time is 0.0, anything else throws csm::Error::BOUNDS
      if (line[i] != 0.0) {
        st = USGSCSM_B200_ST_ERROR_OTHER;
        pos[0] = pos[1] = pos[2] = vel[0] = vel[1] = vel[2] = nan("");
      }
      storeSensorState(o, i, line[i], pos, vel, st);
      continue;
    }
    const double l = line[i], sm = samp[i];
    if ((o.vx || o.vy || o.vz) && (sm < 0.0 || sm > S.nSamples || l < 0.0 || l > S.nLines)) {
      st = USGSCSM_B200_ST_ERROR_OTHER;
      vel[0] = vel[1] = vel[2] = nan("");
    }
    storeSensorState(o, i, 0.0, pos, vel, st);
  }
}

}  // namespace

cudaError_t launchFrSensorState(const FrameState &S, long n, const double *line,
                                const double *samp, const SensorStateOut &o, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  const long blocks = (n + 255) / 256;
  frSensorStateKernel<<<(int)(blocks < 148 * 8 ? blocks : 148 * 8), 256, 0, stream>>>(S, n, line, samp, o);
  countLaunch();
  return cudaGetLastError();
}

// The compile-time variants: the distortion types the frame ISDs carry most (radial, transverse,
// none = radial with zero coefficients is still "radial") without jitter tables; anything else
// runs the generic kernel.  USGSCSM_B200_FR_SPECIALISE=0 forces the generic kernel.
static bool frSpecialise() {
  static const bool on = [] {
    const char *v = getenv("USGSCSM_B200_FR_SPECIALISE");
    return !(v && *v == '0');
  }();
  return on;
}
#define UB_FR_DISPATCH(LAUNCH, VEC)                                              \
  do {                                                                           \
    if (!frSpecialise() || S.numLineTimes) LAUNCH(-1, true, 1);                  \
    else if (S.fp.distortionType == DIST_RADIAL) LAUNCH(DIST_RADIAL, false, VEC);         \
    else if (S.fp.distortionType == DIST_TRANSVERSE && transverseCoeffsPlain(S.fp)) LAUNCH(DIST_TRANSVERSE, false, VEC); \
    else LAUNCH(-1, false, 1);                                                   \
  } while (0)

// every non-null array on a 16-byte boundary (status: 8): two points per 128-bit access
static bool frPairable(std::initializer_list<const void *> doubles, const void *status) {
  static const bool on = [] {
    const char *v = getenv("USGSCSM_B200_FR_PAIRS");
    return !(v && *v == '0');
  }();
  if (!on || (reinterpret_cast<uintptr_t>(status) & 7)) return false;
  for (const void *p : doubles)
    if (reinterpret_cast<uintptr_t>(p) & 15) return false;
  return true;
}

cudaError_t launchFrGroundToImage(const FrameState &S, int smCount, long n, const double *x,
                                  const double *y, const double *z, double prec, double *line,
                                  double *samp, double *achieved, uint32_t *status,
                                  cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
#define UB_FR_G2I(DIST, JITTER, VEC)                                                                  \
  frGroundToImageKernel<DIST, JITTER, VEC>                                                            \
      <<<frBlocks(frGroundToImageKernel<DIST, JITTER, VEC>, (n + VEC - 1) / VEC, smCount), kThreads,  \
         0, stream>>>(S, n, x, y, z, prec, line, samp, achieved, status)
  if (frPairable({x, y, z, line, samp, achieved}, status))
    UB_FR_DISPATCH(UB_FR_G2I, 2);
  else
    UB_FR_DISPATCH(UB_FR_G2I, 1);
#undef UB_FR_G2I
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchFrImageToGround(const FrameState &S, int smCount, long n, const double *line,
                                  const double *samp, const GridSpec *grid, const double *height,
                                  int heightStride, double prec, double *x, double *y, double *z,
                                  double *achieved, uint32_t *status, cudaStream_t stream) {
  (void)prec;  // the frame model's imageToGround is closed form: desiredPrecision is unused
  if (n <= 0) return cudaSuccess;
  GridSpec g = grid ? *grid : GridSpec{0, 0, 0, 0, 1, 0};
#define UB_FR_I2G_OF(GRID, DIST, JITTER, VEC)                                                         \
  frImageToGroundKernel<GRID, DIST, JITTER, VEC>                                                      \
      <<<frBlocks(frImageToGroundKernel<GRID, DIST, JITTER, VEC>, (n + VEC - 1) / VEC, smCount),      \
         kThreads, 0, stream>>>(S, n, line, samp, g, height, heightStride, x, y, z, achieved, status)
#define UB_FR_I2G_GRID(DIST, JITTER, VEC) UB_FR_I2G_OF(true, DIST, JITTER, VEC)
#define UB_FR_I2G_ARRAY(DIST, JITTER, VEC) UB_FR_I2G_OF(false, DIST, JITTER, VEC)
  const bool pairs = frPairable({x, y, z, achieved, heightStride ? height : nullptr,
                                 grid ? nullptr : line, grid ? nullptr : samp}, status);
  if (grid) {
    // rows of an even number of samples
    if (pairs && g.nSamps % 2 == 0)
      UB_FR_DISPATCH(UB_FR_I2G_GRID, 2);
    else
      UB_FR_DISPATCH(UB_FR_I2G_GRID, 1);
  } else if (pairs) {
    UB_FR_DISPATCH(UB_FR_I2G_ARRAY, 2);
  } else {
    UB_FR_DISPATCH(UB_FR_I2G_ARRAY, 1);
  }
#undef UB_FR_I2G_ARRAY
#undef UB_FR_I2G_GRID
#undef UB_FR_I2G_OF
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchFrRemoteLocus(const FrameState &S, int smCount, long n, const double *line,
                                const double *samp, const LocusOut &o, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  frRemoteLocusKernel<<<frBlocks(frRemoteLocusKernel, n, smCount), kThreads, 0, stream>>>(S, n, line, samp, o);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchGroundPartialsFr(const FrameState &S, int smCount, long n, const double *x,
                                   const double *y, const double *z, double *partials,
                                   uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  frGroundPartialsKernel<<<frBlocks(frGroundPartialsKernel, n, smCount), kThreads, 0, stream>>>(S, n, x, y, z, partials, status);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchFrProximateLocus(const FrameState &S, int smCount, long n, const double *line,
                                   const double *samp, const double *x, const double *y,
                                   const double *z, const LocusOut &o, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  frProximateLocusKernel<<<frBlocks(frProximateLocusKernel, n, smCount), kThreads, 0, stream>>>(S, n, line, samp, x, y, z, o);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchJitter(const FrameState &S, int remove, double tolerance, int maxIts, long n,
                         const double *line, const double *sample, double *outLine,
                         double *outSample, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  jitterKernel<<<(int)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184), 256, 0, stream>>>(
      S, remove, tolerance, maxIts, n, line, sample, outLine, outSample);
  countLaunch();
  return cudaGetLastError();
}

}  // namespace ub200
