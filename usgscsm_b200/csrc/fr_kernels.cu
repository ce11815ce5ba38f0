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
#include "kernels.h"
#include "device_math.cuh"

namespace ub200 {

namespace {

constexpr int kThreads = 256;

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

UB_DEV void frGroundToImagePoint(const FrameState &S, double x, double y, double z,
                                 double desiredPrecision, double &outLine, double &outSamp,
                                 uint32_t &status) {
  const double *m = S.mRaw;  // m[r][c] = m[3 * r + c]
  double xo = x - S.params[0];
  double yo = y - S.params[1];
  double zo = z - S.params[2];
  double f = S.fp.focalLength;
  double denom = m[2] * xo + m[5] * yo + m[8] * zo;
  double undistortedx = (f * (m[0] * xo + m[3] * yo + m[6] * zo) / denom);
  double undistortedy = (f * (m[1] * xo + m[4] * yo + m[7] * zo) / denom);
  double distortedX, distortedY;
  applyDistortion(undistortedx, undistortedy, distortedX, distortedY, S.fp, 1.0E-6, status);
  double line, sample;
  computePixel(distortedX, distortedY, S.fp, line, sample);
  if (S.numLineTimes) {
    double jitteredLine, jitteredSample;
    addJitter(S, line, sample, desiredPrecision, 20, jitteredLine, jitteredSample);
    line = jitteredLine;
    sample = jitteredSample;
  }
  outLine = line;
  outSamp = sample;
}

UB_DEV void frImageToGroundPoint(const FrameState &S, double line, double sample, double height,
                                 double &x, double &y, double &z, uint32_t &status) {
  const double *m = S.mNorm;
  if (S.numLineTimes) {
    double dejitteredLine, dejitteredSample;
    removeJitter(S, line, sample, dejitteredLine, dejitteredSample);
    line = dejitteredLine;
    sample = dejitteredSample;
  }
  double x_camera, y_camera;
  distortedFocalPlaneCoordinates(line, sample, S.fp, x_camera, y_camera);
  double undistortedX, undistortedY;
  removeDistortion(x_camera, y_camera, undistortedX, undistortedY, S.fp, 1.0E-6, status);
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

__global__ void __launch_bounds__(kThreads)
    frGroundToImageKernel(const __grid_constant__ FrameState S, long n, const double *__restrict__ x,
                          const double *__restrict__ y, const double *__restrict__ z, double prec,
                          double *__restrict__ line, double *__restrict__ samp,
                          double *__restrict__ achieved, uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double gx = __ldcs(x + i), gy = __ldcs(y + i), gz = __ldcs(z + i);
    double l, s;
    uint32_t st = 0;
    frGroundToImagePoint(S, gx, gy, gz, prec, l, s, st);
    if (!(isfinite(l) && isfinite(s))) st |= USGSCSM_B200_ST_NONFINITE;
    // getValidImageRange FR:600-608: (startingDetectorLine, startingDetectorSample) .. (nLines, nSamples)
    if (!(l >= S.fp.startingDetectorLine && l <= (double)S.nLines &&
          s >= S.fp.startingDetectorSample && s <= (double)S.nSamples))
      st |= USGSCSM_B200_ST_OUT_OF_IMAGE;
    __stcs(line + i, l);
    __stcs(samp + i, s);
    if (achieved) __stcs(achieved + i, nan(""));  // the frame model never writes achievedPrecision
    if (status) __stcs(status + i, st);
  }
}

template <bool GRID>
__global__ void __launch_bounds__(kThreads)
    frImageToGroundKernel(const __grid_constant__ FrameState S, long n, const double *__restrict__ line,
                          const double *__restrict__ samp, GridSpec g,
                          const double *__restrict__ height, int heightStride,
                          double *__restrict__ x, double *__restrict__ y, double *__restrict__ z,
                          double *__restrict__ achieved, uint32_t *__restrict__ status) {
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
    double px, py, pz;
    uint32_t st = 0;
    frImageToGroundPoint(S, l, s, h, px, py, pz, st);
    __stcs(x + i, px);
    __stcs(y + i, py);
    __stcs(z + i, pz);
    if (achieved) __stcs(achieved + i, nan(""));
    if (status) __stcs(status + i, st);
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

// One wave: as many blocks as are resident at once for THIS kernel (its register count decides:
// 3 x 256 threads per SM at 80 registers, 4 at 64), so the grid-stride loop splits the points
// evenly and no block waits for a second wave.
template <class K>
int frBlocks(K kernel, long n, int smCount) {
  static const int perSm = [&] {
    int b = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, kernel, kThreads, 0) != cudaSuccess || b < 1)
      b = 4;
    return b;
  }();
  long blocksNeeded = (n + kThreads - 1) / kThreads;
  long persistent = (long)smCount * perSm;
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

cudaError_t launchFrGroundToImage(const FrameState &S, int smCount, long n, const double *x,
                                  const double *y, const double *z, double prec, double *line,
                                  double *samp, double *achieved, uint32_t *status,
                                  cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  frGroundToImageKernel<<<frBlocks(frGroundToImageKernel, n, smCount), kThreads, 0, stream>>>(S, n, x, y, z, prec, line,
                                                                      samp, achieved, status);
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
  if (grid)
    frImageToGroundKernel<true><<<frBlocks(frImageToGroundKernel<true>, n, smCount), kThreads, 0, stream>>>(
        S, n, line, samp, g, height, heightStride, x, y, z, achieved, status);
  else
    frImageToGroundKernel<false><<<frBlocks(frImageToGroundKernel<false>, n, smCount), kThreads, 0, stream>>>(
        S, n, line, samp, g, height, heightStride, x, y, z, achieved, status);
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
