// illumination.cu -- getIlluminationDirection(const csm::EcefCoord&) for arrays
// (LS:1878-1899 + getSunPosition LS:3165-3193, PF:1477-1498 + PF:2466-2494, SAR:1095-1102 +
// SAR:1974-2004, FR:643-652).
//
// Line-scan, push-frame, SAR: the ground point is projected (groundToImage with its default
// precision 0.001 and no warning list), the image time of that point selects the sun position
// -- Lagrange-8 over the m_sunPosition table, or m_sunPosition + t * m_sunVelocity when the
// state holds one position, or the position alone -- and the direction sun -> ground is
// normalised (component / |v| for LS / PF, (1 / |v|) * component for SAR, as the reference
// writes them).  Frame: ground - m_sunPosition, not normalised, no projection.
#include <vector>

#include "capi_common.h"
#include "kernels.h"
#include "state_reader.h"

namespace ub200 {

namespace {

constexpr uint32_t kErrBits = USGSCSM_B200_ST_ERROR_ALGORITHM | USGSCSM_B200_ST_ERROR_UNKNOWN |
                              USGSCSM_B200_ST_ERROR_OTHER;

struct SunModel {
  int mode;       // 0: constant position, 1: position + time * velocity, 2: interpolated (sunXyz)
  int normalise;  // 0: none (frame), 1: v / |v| (LS, PF), 2: (1 / |v|) * v (SAR)
  double pos[3], vel[3];
};

__global__ void illuminationKernel(long n, SunModel sun, const double *__restrict__ x,
                                   const double *__restrict__ y, const double *__restrict__ z,
                                   const double *__restrict__ time,
                                   const double *__restrict__ sunXyz,
                                   const uint32_t *__restrict__ stProjection,
                                   double *__restrict__ dx, double *__restrict__ dy,
                                   double *__restrict__ dz, uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double s[3];
    if (sun.mode == 2) {
      s[0] = sunXyz[3 * i];
      s[1] = sunXyz[3 * i + 1];
      s[2] = sunXyz[3 * i + 2];
    } else if (sun.mode == 1) {
      const double t = time[i];
      s[0] = t * sun.vel[0] + sun.pos[0];
      s[1] = t * sun.vel[1] + sun.pos[1];
      s[2] = t * sun.vel[2] + sun.pos[2];
    } else {
      s[0] = sun.pos[0];
      s[1] = sun.pos[1];
      s[2] = sun.pos[2];
    }
    double v[3] = {x[i] - s[0], y[i] - s[1], z[i] - s[2]};
    if (sun.normalise) {
      const double scale = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
      if (sun.normalise == 1) {
        v[0] /= scale;
        v[1] /= scale;
        v[2] /= scale;
      } else {
        const double r = 1.0 / scale;
        v[0] = r * v[0];
        v[1] = r * v[1];
        v[2] = r * v[2];
      }
    }
    uint32_t st = stProjection ? (stProjection[i] & kErrBits) : 0u;
    if (st) v[0] = v[1] = v[2] = nan("");  // groundToImage would have thrown
    if (!(isfinite(v[0]) && isfinite(v[1]) && isfinite(v[2]))) st |= USGSCSM_B200_ST_NONFINITE;
    dx[i] = v[0];
    dy[i] = v[1];
    dz[i] = v[2];
    if (status) status[i] = st;
  }
}

}  // namespace

// m_sunPosition / m_sunVelocity of the state; the interpolation table when there is more than
// one position (t0 / dt: the ephemeris grid for LS and PF, LS:3174-3175; the image duration
// over the node count for SAR, SAR:1983-1986).
void loadSun(Model &m, const StateReader &st) {
  m.sunPosition = st.getVecDouble("m_sunPosition", std::vector<double>());
  m.sunVelocity = st.getVecDouble("m_sunVelocity", std::vector<double>());
  m.hasSunTable = false;
  const int nodes = (int)(m.sunPosition.size() / 3);
  if (m.kind == USGSCSM_B200_MODEL_FRAME || nodes <= 1) return;
  double t0, dt;
  if (m.kind == USGSCSM_B200_MODEL_SAR) {
    t0 = m.sar.positions.t0;
    const double endTime = t0 + m.sar.nLines * m.sar.exposureDuration;
    dt = (endTime - t0) / nodes;
  } else {
    t0 = m.ls.positions.t0;
    dt = m.ls.positions.dt;
  }
  m.sunTable = UniformTable{m.upload(m.sunPosition), nodes, 8, t0, dt};
  fillTableSteps({&m.sunTable});
  std::vector<Divisor *> ds;
  appendLagrangeDivisors(m.sunLag, ds);
  fillDivisors(ds);
  m.hasSunTable = true;
}

int illuminationDev(usgscsm_b200_model *m, long n, const double *x, const double *y,
                    const double *z, double *dx, double *dy, double *dz, uint32_t *status,
                    cudaStream_t stream) {
  if (m->sunPosition.size() < 3)
    throw std::invalid_argument("the model state holds no m_sunPosition");
  SunModel sun{};
  sun.mode = m->hasSunTable ? 2 : (m->sunVelocity.size() / 3 >= 1 ? 1 : 0);
  sun.normalise = m->kind == USGSCSM_B200_MODEL_FRAME ? 0 : m->kind == USGSCSM_B200_MODEL_SAR ? 2 : 1;
  for (int k = 0; k < 3; k++) {
    sun.pos[k] = m->sunPosition[k];
    sun.vel[k] = m->sunVelocity.size() >= 3 ? m->sunVelocity[k] : 0.0;
  }
  const long blocks = (n + 255) / 256;
  const int grid = (int)(blocks < 148 * 8 ? blocks : 148 * 8);
  if (m->kind == USGSCSM_B200_MODEL_FRAME) {
    sun.mode = 0;  // FR:643-652 reads m_sunPosition[0..2] only
    illuminationKernel<<<grid, 256, 0, stream>>>(n, sun, x, y, z, nullptr, nullptr, nullptr, dx, dy,
                                                 dz, status);
    countLaunch();
    UB_CUDA(cudaGetLastError());
    return USGSCSM_B200_OK;
  }
  double *scratch = nullptr;
  uint32_t *stProj = nullptr;
  UB_CUDA(cudaMallocAsync(&scratch, (size_t)n * 6 * sizeof(double), stream));
  UB_CUDA(cudaMallocAsync(&stProj, (size_t)n * sizeof(uint32_t), stream));
  double *line = scratch, *samp = scratch + n, *time = scratch + 2 * n, *sunXyz = scratch + 3 * n;
  cudaError_t e = launchGroundToImage(*m, n, x, y, z, 0.001, line, samp, nullptr, stProj, stream);
  if (e == cudaSuccess) {
    SensorStateOut o{time, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    switch (m->kind) {
      case USGSCSM_B200_MODEL_LINE_SCANNER: e = launchLsSensorState(m->ls, n, line, o, stream); break;
      case USGSCSM_B200_MODEL_PUSH_FRAME: e = launchPfSensorState(m->ls, n, line, o, stream); break;
      default: e = launchSarSensorState(m->sar, n, line, o, stream); break;
    }
  }
  if (e == cudaSuccess && sun.mode == 2)
    e = launchLagrange(m->sunTable, m->sunLag, 3, n, time, sunXyz, stream);
  if (e == cudaSuccess) {
    illuminationKernel<<<grid, 256, 0, stream>>>(n, sun, x, y, z, time, sunXyz, stProj, dx, dy, dz,
                                                 status);
    countLaunch();
    e = cudaGetLastError();
  }
  cudaFreeAsync(scratch, stream);
  cudaFreeAsync(stProj, stream);
  UB_CUDA(e);
  return USGSCSM_B200_OK;
}

}  // namespace ub200

using namespace ub200;

extern "C" int usgscsm_b200_illumination_direction_dev(usgscsm_b200_model *m, int64_t n,
                                                       const double *x, const double *y,
                                                       const double *z, double *dx, double *dy,
                                                       double *dz, uint32_t *status, void *stream) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!x || !y || !z || !dx || !dy || !dz) return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    return illuminationDev(m, (long)n, x, y, z, dx, dy, dz, status, static_cast<cudaStream_t>(stream));
  });
}
