// geo_kernels.cu -- kernels of the georeferenced-grid support (geogrid.h, geo_device.cuh): DEM
// window -> ground points, raster pixel <-> body-fixed coordinates (what
// UsgsAstroProjectedSensorModel wraps around a camera, src/UsgsAstroProjectedSensorModel.cpp:
// 289-365), and the height update of the DEM-iterated ground intersection.
#include "geo_device.cuh"
#include "geo_kernels.h"
#include "kernels.h"

namespace ub200 {

namespace {

constexpr int kGeoThreads = 256;

int geoBlocks(long n, int smCount) {
  long need = (n + kGeoThreads - 1) / kGeoThreads;
  long cap = (long)smCount * 8;
  return (int)(need < cap ? (need < 1 ? 1 : need) : cap);
}

__global__ void __launch_bounds__(kGeoThreads)
    demGroundPointsKernel(const __grid_constant__ DemWindow w, long n, double *__restrict__ x,
                          double *__restrict__ y, double *__restrict__ z,
                          uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double gx, gy, gz;
    uint32_t st = 0;
    if (!demGroundPoint(w, i, gx, gy, gz)) st |= USGSCSM_B200_ST_DEM_NODATA;
    if (!finite3(gx, gy, gz)) st |= USGSCSM_B200_ST_NONFINITE;
    __stcs(x + i, gx);
    __stcs(y + i, gy);
    __stcs(z + i, gz);
    if (status) __stcs(status + i, st);
  }
}

// After a model's own groundToImage kernel ran on generated ground points: a nodata sample's
// outputs become NaN / DEM_NODATA | NONFINITE whatever the search made of the NaN input.
__global__ void __launch_bounds__(kGeoThreads)
    demNodataFixupKernel(const __grid_constant__ DemWindow w, long n, double *__restrict__ line,
                         double *__restrict__ samp, double *__restrict__ achieved,
                         uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const long r = i / w.nCols, c = i - r * w.nCols;
    double raw;
    if (demRaw(w.samples, w.sampleType, r * w.rowStride + c, w.hasNodata, w.nodata, raw)) continue;
    line[i] = samp[i] = nan("");
    if (achieved) achieved[i] = nan("");
    if (status) status[i] = USGSCSM_B200_ST_DEM_NODATA | USGSCSM_B200_ST_NONFINITE;
  }
}

// UsgsAstroProjectedSensorModel::imageToGround, src/UsgsAstroProjectedSensorModel.cpp:329-365
__global__ void __launch_bounds__(kGeoThreads)
    gridImageToGroundKernel(const __grid_constant__ GeoGrid g, long n, const double *__restrict__ line,
                            const double *__restrict__ samp, const double *__restrict__ height,
                            int heightStride, double *__restrict__ x, double *__restrict__ y,
                            double *__restrict__ z, uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double h = height ? (heightStride ? __ldcs(height + i) : __ldg(height)) : 0.0;
    double mx, my, lam, phi, gx, gy, gz;
    geoPixelToMeter(g, __ldcs(line + i), __ldcs(samp + i), mx, my);
    if (geoProjInverse(g, mx, my, lam, phi)) geoCartForward(g, lam, phi, h, gx, gy, gz);
    else gx = gy = gz = nan("");
    __stcs(x + i, gx);
    __stcs(y + i, gy);
    __stcs(z + i, gz);
    if (status) __stcs(status + i, finite3(gx, gy, gz) ? 0u : (uint32_t)USGSCSM_B200_ST_NONFINITE);
  }
}

// UsgsAstroProjectedSensorModel::groundToImage, src/UsgsAstroProjectedSensorModel.cpp:289-321
__global__ void __launch_bounds__(kGeoThreads)
    gridGroundToImageKernel(const __grid_constant__ GeoGrid g, long n, const double *__restrict__ x,
                            const double *__restrict__ y, const double *__restrict__ z,
                            double *__restrict__ line, double *__restrict__ samp,
                            double *__restrict__ height, uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double lam, phi, h, mx, my, l, s;
    geoCartInverse(g, __ldcs(x + i), __ldcs(y + i), __ldcs(z + i), lam, phi, h);
    geoProjForward(g, lam, phi, mx, my);
    geoMeterToPixel(g, mx, my, l, s);
    __stcs(line + i, l);
    __stcs(samp + i, s);
    if (height) __stcs(height + i, h);
    if (status) __stcs(status + i, (isfinite(l) && isfinite(s)) ? 0u : (uint32_t)USGSCSM_B200_ST_NONFINITE);
  }
}

// Bilinear DEM height at raster coordinates (line, samp); pixel centres at +0.5.
// 0 = ok, otherwise the status bit to raise.
UB_DEV uint32_t demBilinear(const DemRaster &d, double line, double samp, double &h) {
  const double u = samp - 0.5, v = line - 0.5;
  if (!(u >= 0.0 && v >= 0.0 && u <= (double)(d.nCols - 1) && v <= (double)(d.nRows - 1)))
    return USGSCSM_B200_ST_DEM_OUTSIDE;
  const double fc = floor(u), fr = floor(v);
  long c0 = (long)fc, r0 = (long)fr;
  long c1 = c0 + 1 < d.nCols ? c0 + 1 : c0;
  long r1 = r0 + 1 < d.nRows ? r0 + 1 : r0;
  const double fu = u - fc, fv = v - fr;
  double h00, h01, h10, h11;
  bool ok = demRaw(d.samples, d.sampleType, r0 * d.nCols + c0, d.hasNodata, d.nodata, h00);
  ok = demRaw(d.samples, d.sampleType, r0 * d.nCols + c1, d.hasNodata, d.nodata, h01) && ok;
  ok = demRaw(d.samples, d.sampleType, r1 * d.nCols + c0, d.hasNodata, d.nodata, h10) && ok;
  ok = demRaw(d.samples, d.sampleType, r1 * d.nCols + c1, d.hasNodata, d.nodata, h11) && ok;
  if (!ok) return USGSCSM_B200_ST_DEM_NODATA;
  h00 = h00 * d.scale + d.offset;
  h01 = h01 * d.scale + d.offset;
  h10 = h10 * d.scale + d.offset;
  h11 = h11 * d.scale + d.offset;
  const double top = h00 * (1.0 - fu) + h01 * fu;
  const double bot = h10 * (1.0 - fu) + h11 * fu;
  h = top * (1.0 - fv) + bot * fv;
  return 0;
}

constexpr uint32_t kDemDone = 1u << 30;  // internal: the point's height has settled (or failed)

// One step of the DEM fixed point after imageToGround ran at heights h[]: status[] holds the
// bits of that imageToGround; points already done keep everything.
__global__ void __launch_bounds__(kGeoThreads)
    demHeightUpdateKernel(const __grid_constant__ DemRaster d, long n, const double *__restrict__ x,
                          const double *__restrict__ y, const double *__restrict__ z,
                          double *__restrict__ h, const uint32_t *__restrict__ i2gStatus,
                          uint32_t *__restrict__ state, double tolerance, int last) {
  const long stride = (long)gridDim.x * blockDim.x;
  const uint32_t kErr = USGSCSM_B200_ST_ERROR_ALGORITHM | USGSCSM_B200_ST_ERROR_UNKNOWN |
                        USGSCSM_B200_ST_ERROR_OTHER | USGSCSM_B200_ST_NONFINITE;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t st = state[i];
    if (st & kDemDone) {
      if (last) state[i] = st & ~kDemDone;
      continue;
    }
    st = i2gStatus[i] & ~USGSCSM_B200_ST_EVALS_MASK;
    if (st & kErr) {
      st |= kDemDone;
    } else {
      double lam, phi, hp, mx, my, l, s, hn;
      geoCartInverse(d.grid, x[i], y[i], z[i], lam, phi, hp);
      geoProjForward(d.grid, lam, phi, mx, my);
      geoMeterToPixel(d.grid, mx, my, l, s);
      const uint32_t bad = demBilinear(d, l, s, hn);
      if (bad) {
        st |= bad | kDemDone;
      } else if (fabs(hn - h[i]) <= tolerance) {
        st |= kDemDone;
      } else if (last) {
        st |= USGSCSM_B200_ST_ITERATION_CAP;
      } else {
        h[i] = hn;
      }
    }
    if (last) st &= ~kDemDone;
    state[i] = st;
  }
}

__global__ void fillKernel(long n, double v, double *__restrict__ a, uint32_t *__restrict__ st) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    a[i] = v;
    st[i] = 0;
  }
}

}  // namespace

cudaError_t launchDemGroundPoints(const DemWindow &w, int smCount, long n, double *x, double *y,
                                  double *z, uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  demGroundPointsKernel<<<geoBlocks(n, smCount), kGeoThreads, 0, stream>>>(w, n, x, y, z, status);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchDemNodataFixup(const DemWindow &w, int smCount, long n, double *line, double *samp,
                                 double *achieved, uint32_t *status, cudaStream_t stream) {
  if (n <= 0 || !w.hasNodata) return cudaSuccess;
  demNodataFixupKernel<<<geoBlocks(n, smCount), kGeoThreads, 0, stream>>>(w, n, line, samp, achieved,
                                                                        status);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchGridImageToGround(const GeoGrid &g, int smCount, long n, const double *line,
                                    const double *samp, const double *height, int heightStride,
                                    double *x, double *y, double *z, uint32_t *status,
                                    cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  gridImageToGroundKernel<<<geoBlocks(n, smCount), kGeoThreads, 0, stream>>>(
      g, n, line, samp, height, heightStride, x, y, z, status);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchGridGroundToImage(const GeoGrid &g, int smCount, long n, const double *x,
                                    const double *y, const double *z, double *line, double *samp,
                                    double *height, uint32_t *status, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  gridGroundToImageKernel<<<geoBlocks(n, smCount), kGeoThreads, 0, stream>>>(g, n, x, y, z, line,
                                                                            samp, height, status);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchDemHeightUpdate(const DemRaster &d, int smCount, long n, const double *x,
                                  const double *y, const double *z, double *h,
                                  const uint32_t *i2gStatus, uint32_t *state, double tolerance,
                                  int last, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  demHeightUpdateKernel<<<geoBlocks(n, smCount), kGeoThreads, 0, stream>>>(d, n, x, y, z, h, i2gStatus,
                                                                          state, tolerance, last);
  countLaunch();
  return cudaGetLastError();
}

cudaError_t launchFill(int smCount, long n, double v, double *a, uint32_t *st, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  fillKernel<<<geoBlocks(n, smCount), kGeoThreads, 0, stream>>>(n, v, a, st);
  countLaunch();
  return cudaGetLastError();
}

}  // namespace ub200
