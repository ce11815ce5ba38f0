// capi_geo.cu -- C-ABI of the georeferenced-grid entry points (include/usgscsm_b200.h, "the
// callers either side of the path"): raster pixel <-> body-fixed coordinates, DEM window ->
// ground points, groundToImage over a DEM window with the ground points generated in the
// kernel, and the DEM-iterated ground intersection.
#include <cstring>
#include <memory>
#include <vector>

#include "capi_common.h"
#include "geo_kernels.h"
#include "geogrid.h"
#include "pipeline.cuh"

namespace ub200 {
usgscsm_b200_geogrid geoGridCFromProjString(const std::string &proj, const double gt[6],
                                            double cartA, double cartB);  // geogrid.cpp

namespace {

int sampleBytes(int type) {
  switch (type) {
    case USGSCSM_B200_DEM_F64: return 8;
    case USGSCSM_B200_DEM_F32: return 4;
    case USGSCSM_B200_DEM_I16: return 2;
    default: return 0;
  }
}

int smCountOf(int device) {
  int n = 148;
  cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
  return n;
}

int checkDevice(int device) {
  if (device < 0 || device >= deviceCount())
    return fail(USGSCSM_B200_ENODEVICE, "no usable CUDA device with that index");
  return USGSCSM_B200_OK;
}

int checkDem(const usgscsm_b200_dem *dem, int64_t row0, int64_t col0, int64_t nRows, int64_t nCols) {
  if (!dem || !dem->samples) return fail(USGSCSM_B200_EINVAL, "null DEM");
  if (!sampleBytes(dem->sample_type)) return fail(USGSCSM_B200_EINVAL, "unknown DEM sample type");
  if (row0 < 0 || col0 < 0 || nRows < 0 || nCols < 0 || row0 + nRows > dem->n_rows ||
      col0 + nCols > dem->n_cols)
    return fail(USGSCSM_B200_EINVAL, "window outside the DEM raster");
  return USGSCSM_B200_OK;
}

// Tables of a separable grid window on the device (stream-ordered scratch).  The host vectors
// are uploaded with a plain cudaMemcpyAsync from pageable memory, which returns once the
// bytes are staged, so they may be dropped on return.
struct WindowTables {
  double *dev = nullptr;
  cudaStream_t stream = nullptr;
  const double *rowTab = nullptr, *colTab = nullptr;
  // With a model: the tables live in the model's small cache (kept until evicted or the model is
  // freed); a hit costs nothing.
  void build(Model &m, const GeoGrid &g, int64_t row0, int64_t col0, int64_t nRows, int64_t nCols) {
    if (!geoGridSeparable(g)) return;
    std::string key(reinterpret_cast<const char *>(&g), sizeof(g));
    const int64_t win[4] = {row0, col0, nRows, nCols};
    key.append(reinterpret_cast<const char *>(win), sizeof(win));
    std::lock_guard<std::mutex> lock(m.geoTablesMutex);
    for (Model::GeoTables &t : m.geoTables)
      if (t.key == key) {
        rowTab = t.dev;
        colTab = t.dev + t.rowDoubles;
        return;
      }
    std::vector<double> rows, cols;
    geoGridTables(g, (double)row0 + 0.5, 1.0, nRows, (double)col0 + 0.5, 1.0, nCols, rows, cols);
    Model::GeoTables t;
    t.key = key;
    t.rowDoubles = rows.size();
    UB_CUDA(cudaMalloc(&t.dev, (rows.size() + cols.size()) * sizeof(double)));
    UB_CUDA(cudaMemcpy(t.dev, rows.data(), rows.size() * sizeof(double), cudaMemcpyHostToDevice));
    UB_CUDA(cudaMemcpy(t.dev + rows.size(), cols.data(), cols.size() * sizeof(double),
                       cudaMemcpyHostToDevice));
    if (m.geoTables.size() >= 4) {  // evict the oldest; cudaFree waits for work that may read it
      cudaFree(m.geoTables.front().dev);
      m.geoTables.erase(m.geoTables.begin());
    }
    m.geoTables.push_back(t);
    rowTab = t.dev;
    colTab = t.dev + t.rowDoubles;
  }
  void build(const GeoGrid &g, int64_t row0, int64_t col0, int64_t nRows, int64_t nCols,
             cudaStream_t s) {
    if (!geoGridSeparable(g)) return;
    std::vector<double> rows, cols;
    geoGridTables(g, (double)row0 + 0.5, 1.0, nRows, (double)col0 + 0.5, 1.0, nCols, rows, cols);
    stream = s;
    UB_CUDA(cudaMallocAsync(&dev, (rows.size() + cols.size()) * sizeof(double), s));
    UB_CUDA(cudaMemcpyAsync(dev, rows.data(), rows.size() * sizeof(double), cudaMemcpyHostToDevice, s));
    UB_CUDA(cudaMemcpyAsync(dev + rows.size(), cols.data(), cols.size() * sizeof(double),
                            cudaMemcpyHostToDevice, s));
    rowTab = dev;
    colTab = dev + rows.size();
  }
  ~WindowTables() {
    if (dev) cudaFreeAsync(dev, stream);
  }
};

// The window [row0, row0 + nRows) x [col0, col0 + nCols) of a raster whose samples are at
// `samples` with `rowStride` samples per row; tables (if any) cover the window from its row
// `tabRow0` on.
DemWindow makeWindow(const usgscsm_b200_dem &dem, const GeoGrid &g, const void *samples,
                     long rowStride, int64_t row0, int64_t col0, int64_t nCols,
                     const WindowTables &t, int64_t tabRow0) {
  DemWindow w{};
  w.samples = samples;
  w.sampleType = dem.sample_type;
  w.hasNodata = dem.has_nodata;
  w.nodata = dem.nodata;
  w.scale = dem.scale;
  w.offset = dem.offset;
  w.rowStride = rowStride;
  w.nCols = (long)nCols;
  w.rowTab = t.rowTab ? t.rowTab + 4 * tabRow0 : nullptr;
  w.colTab = t.colTab;
  w.grid = g;
  w.line0 = (double)row0 + 0.5;
  w.samp0 = (double)col0 + 0.5;
  return w;
}

// groundToImage over a window whose samples are on the device.
cudaError_t launchGroundToImageDem(Model &m, const DemWindow &w, long n, double prec, double *line,
                                   double *samp, double *achieved, uint32_t *status,
                                   cudaStream_t stream) {
  if (m.kind == USGSCSM_B200_MODEL_LINE_SCANNER)
    return launchLsGroundToImageDem(m.ls, m.smCount, n, w, prec, line, samp, achieved, status, stream);
  double *xyz = nullptr;
  cudaError_t e = cudaMallocAsync(&xyz, (size_t)n * 3 * sizeof(double), stream);
  if (e != cudaSuccess) return e;
  e = launchDemGroundPoints(w, m.smCount, n, xyz, xyz + n, xyz + 2 * n, nullptr, stream);
  if (e == cudaSuccess)
    e = launchGroundToImage(m, n, xyz, xyz + n, xyz + 2 * n, prec, line, samp, achieved, status, stream);
  if (e == cudaSuccess) e = launchDemNodataFixup(w, m.smCount, n, line, samp, achieved, status, stream);
  cudaError_t e2 = cudaFreeAsync(xyz, stream);
  return e != cudaSuccess ? e : e2;
}

const void *sampleAt(const usgscsm_b200_dem &dem, int64_t row, int64_t col) {
  return static_cast<const char *>(dem.samples) + (row * dem.n_cols + col) * sampleBytes(dem.sample_type);
}

// A model-less pipeline owner for the grid-only host entry points: they reuse runPipelined, which
// keeps its slots in a Model.
usgscsm_b200_model *scratchModel(int device) {
  // never destroyed: their streams and buffers must not be released from a static destructor,
  // after the CUDA runtime has begun to unload
  static std::mutex mu;
  static std::vector<usgscsm_b200_model *> *perDevice = new std::vector<usgscsm_b200_model *>();
  std::lock_guard<std::mutex> lock(mu);
  if ((int)perDevice->size() <= device) perDevice->resize((size_t)device + 1, nullptr);
  if (!(*perDevice)[(size_t)device]) {
    usgscsm_b200_model *m = new usgscsm_b200_model;
    m->device = device;
    m->smCount = smCountOf(device);
    (*perDevice)[(size_t)device] = m;
  }
  return (*perDevice)[(size_t)device];
}

}  // namespace
}  // namespace ub200

using namespace ub200;

extern "C" {

int usgscsm_b200_geogrid_from_proj_string(const char *proj_string, const double geotransform[6],
                                          double cart_a, double cart_b, usgscsm_b200_geogrid *out) {
  if (!proj_string || !geotransform || !out) return fail(USGSCSM_B200_EINVAL, "null argument");
  try {
    *out = geoGridCFromProjString(proj_string, geotransform, cart_a, cart_b);
    (void)geoGridFromC(*out);  // validates
  } catch (const std::exception &e) {
    lastError() = e.what();
    return USGSCSM_B200_EUNSUPPORTED;
  }
  return USGSCSM_B200_OK;
}

int usgscsm_b200_geogrid_image_to_ground_dev(const usgscsm_b200_geogrid *grid, int device, int64_t n,
                                             const double *line, const double *samp,
                                             const double *height, int height_stride, double *x,
                                             double *y, double *z, uint32_t *status, void *stream) {
  if (!grid || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!line || !samp || !x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  if (int rc = checkDevice(device)) return rc;
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(device));
    const GeoGrid g = geoGridFromC(*grid);
    UB_CUDA(launchGridImageToGround(g, smCountOf(device), (long)n, line, samp, height,
                                    height_stride ? 1 : 0, x, y, z, status,
                                    static_cast<cudaStream_t>(stream)));
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_geogrid_ground_to_image_dev(const usgscsm_b200_geogrid *grid, int device, int64_t n,
                                             const double *x, const double *y, const double *z,
                                             double *line, double *samp, double *height,
                                             uint32_t *status, void *stream) {
  if (!grid || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!x || !y || !z || !line || !samp) return fail(USGSCSM_B200_EINVAL, "null array");
  if (int rc = checkDevice(device)) return rc;
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(device));
    const GeoGrid g = geoGridFromC(*grid);
    UB_CUDA(launchGridGroundToImage(g, smCountOf(device), (long)n, x, y, z, line, samp, height, status,
                                    static_cast<cudaStream_t>(stream)));
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_geogrid_image_to_ground(const usgscsm_b200_geogrid *grid, int device, int64_t n,
                                         const double *line, const double *samp, const double *height,
                                         int height_stride, double *x, double *y, double *z,
                                         uint32_t *status) {
  if (!grid || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!line || !samp || !x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  if (int rc = checkDevice(device)) return rc;
  GeoGrid g;
  int rc = guarded([&]() -> int {
    g = geoGridFromC(*grid);
    return USGSCSM_B200_OK;
  });
  if (rc) return rc;
  const int hs = height_stride ? 1 : 0;
  const double zero = 0.0;
  const double *h = height ? height : &zero;
  const bool bcast = !height || hs == 0;
  usgscsm_b200_model *m = scratchModel(device);
  return runPipelined(m, n, {{line, 1, false}, {samp, 1, false}, {h, 1, bcast}}, {{x, 1}, {y, 1}, {z, 1}},
                      status, 0, [&](const Chunk &k) {
                        return launchGridImageToGround(g, m->smCount, k.count, k.in[0], k.in[1], k.in[2],
                                                       bcast ? 0 : 1, k.out[0], k.out[1], k.out[2],
                                                       k.status, k.stream);
                      });
}

int usgscsm_b200_geogrid_ground_to_image(const usgscsm_b200_geogrid *grid, int device, int64_t n,
                                         const double *x, const double *y, const double *z,
                                         double *line, double *samp, double *height,
                                         uint32_t *status) {
  if (!grid || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!x || !y || !z || !line || !samp) return fail(USGSCSM_B200_EINVAL, "null array");
  if (int rc = checkDevice(device)) return rc;
  GeoGrid g;
  int rc = guarded([&]() -> int {
    g = geoGridFromC(*grid);
    return USGSCSM_B200_OK;
  });
  if (rc) return rc;
  usgscsm_b200_model *m = scratchModel(device);
  return runPipelined(m, n, {{x, 1, false}, {y, 1, false}, {z, 1, false}},
                      {{line, 1}, {samp, 1}, {height, 1}}, status, 0, [&](const Chunk &k) {
                        return launchGridGroundToImage(g, m->smCount, k.count, k.in[0], k.in[1], k.in[2],
                                                       k.out[0], k.out[1], k.out[2], k.status, k.stream);
                      });
}

int usgscsm_b200_dem_ground_points_dev(const usgscsm_b200_dem *dem, int device, int64_t row0,
                                       int64_t col0, int64_t n_rows, int64_t n_cols, double *x,
                                       double *y, double *z, uint32_t *status, void *stream) {
  if (int rc = checkDem(dem, row0, col0, n_rows, n_cols)) return rc;
  if (n_rows == 0 || n_cols == 0) return USGSCSM_B200_OK;
  if (!x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  if (int rc = checkDevice(device)) return rc;
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const GeoGrid g = geoGridFromC(dem->grid);
    WindowTables t;
    t.build(g, row0, col0, n_rows, n_cols, s);
    const DemWindow w = makeWindow(*dem, g, sampleAt(*dem, row0, col0), (long)dem->n_cols, row0, col0,
                                   n_cols, t, 0);
    UB_CUDA(launchDemGroundPoints(w, smCountOf(device), (long)(n_rows * n_cols), x, y, z, status, s));
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_dem_ground_points(const usgscsm_b200_dem *dem, int device, int64_t row0, int64_t col0,
                                   int64_t n_rows, int64_t n_cols, double *x, double *y, double *z,
                                   uint32_t *status) {
  if (int rc = checkDem(dem, row0, col0, n_rows, n_cols)) return rc;
  if (n_rows == 0 || n_cols == 0) return USGSCSM_B200_OK;
  if (!x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  if (int rc = checkDevice(device)) return rc;
  usgscsm_b200_model *m = scratchModel(device);
  GeoGrid g;
  WindowTables t;
  int rc = guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(device));
    g = geoGridFromC(dem->grid);
    t.build(g, row0, col0, n_rows, n_cols, nullptr);
    UB_CUDA(cudaStreamSynchronize(nullptr));
    return USGSCSM_B200_OK;
  });
  if (rc) return rc;
  const int sb = sampleBytes(dem->sample_type);
  return runPipelined(
      m, n_rows * n_cols,
      {HostIn(sampleAt(*dem, row0, col0), sb, (long)n_cols, (long)dem->n_cols * sb)},
      {{x, 1}, {y, 1}, {z, 1}}, status, 0,
      [&](const Chunk &k) {
        const int64_t r = k.offset / n_cols;
        const DemWindow w = makeWindow(*dem, g, k.in[0], (long)n_cols, row0 + r, col0, n_cols, t, r);
        return launchDemGroundPoints(w, m->smCount, k.count, k.out[0], k.out[1], k.out[2], k.status,
                                     k.stream);
      },
      (long)n_cols);
}

int usgscsm_b200_ground_to_image_dem_grid_dev(usgscsm_b200_model *m, const usgscsm_b200_dem *dem,
                                              int64_t row0, int64_t col0, int64_t n_rows,
                                              int64_t n_cols, double prec, double *line, double *samp,
                                              double *achieved, uint32_t *status, void *stream) {
  if (!m) return fail(USGSCSM_B200_EINVAL, "null model");
  if (int rc = checkDem(dem, row0, col0, n_rows, n_cols)) return rc;
  if (n_rows == 0 || n_cols == 0) return USGSCSM_B200_OK;
  if (!line || !samp) return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const GeoGrid g = geoGridFromC(dem->grid);
    WindowTables t;
    t.build(*m, g, row0, col0, n_rows, n_cols);
    const DemWindow w = makeWindow(*dem, g, sampleAt(*dem, row0, col0), (long)dem->n_cols, row0, col0,
                                   n_cols, t, 0);
    UB_CUDA(launchGroundToImageDem(*m, w, (long)(n_rows * n_cols), prec, line, samp, achieved, status, s));
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_ground_to_image_dem_grid(usgscsm_b200_model *m, const usgscsm_b200_dem *dem,
                                          int64_t row0, int64_t col0, int64_t n_rows, int64_t n_cols,
                                          double prec, double *line, double *samp, double *achieved,
                                          uint32_t *status) {
  if (!m) return fail(USGSCSM_B200_EINVAL, "null model");
  if (int rc = checkDem(dem, row0, col0, n_rows, n_cols)) return rc;
  if (n_rows == 0 || n_cols == 0) return USGSCSM_B200_OK;
  if (!line || !samp) return fail(USGSCSM_B200_EINVAL, "null array");
  GeoGrid g;
  WindowTables t;
  int rc = guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    g = geoGridFromC(dem->grid);
    t.build(*m, g, row0, col0, n_rows, n_cols);
    return USGSCSM_B200_OK;
  });
  if (rc) return rc;
  const int sb = sampleBytes(dem->sample_type);
  return runPipelined(
      m, n_rows * n_cols,
      {HostIn(sampleAt(*dem, row0, col0), sb, (long)n_cols, (long)dem->n_cols * sb)},
      {{line, 1}, {samp, 1}, {achieved, 1}}, status, 0,
      [&](const Chunk &k) {
        const int64_t r = k.offset / n_cols;
        const DemWindow w = makeWindow(*dem, g, k.in[0], (long)n_cols, row0 + r, col0, n_cols, t, r);
        return launchGroundToImageDem(*m, w, k.count, prec, k.out[0], k.out[1], k.out[2], k.status,
                                      k.stream);
      },
      (long)n_cols);
}

int usgscsm_b200_image_to_ground_dem_dev(usgscsm_b200_model *m, const usgscsm_b200_dem *dem, int64_t n,
                                         const double *line, const double *samp, double h0,
                                         double tolerance, int max_iterations, double prec, double *x,
                                         double *y, double *z, double *height, uint32_t *status,
                                         void *stream) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (int rc = checkDem(dem, 0, 0, dem ? dem->n_rows : 0, dem ? dem->n_cols : 0)) return rc;
  if (max_iterations < 1 || max_iterations > 64 || !(tolerance >= 0))
    return fail(USGSCSM_B200_EINVAL, "max_iterations must be 1..64 and tolerance >= 0");
  if (n == 0) return USGSCSM_B200_OK;
  if (!line || !samp || !x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    DemRaster d{};
    d.samples = dem->samples;
    d.sampleType = dem->sample_type;
    d.hasNodata = dem->has_nodata;
    d.nodata = dem->nodata;
    d.scale = dem->scale;
    d.offset = dem->offset;
    d.nRows = (long)dem->n_rows;
    d.nCols = (long)dem->n_cols;
    d.grid = geoGridFromC(dem->grid);
    // scratch: current heights (unless the caller wants them), imageToGround status, state
    double *hbuf = height;
    double *scratchH = nullptr;
    uint32_t *scratchSt = nullptr;
    if (!hbuf) {
      UB_CUDA(cudaMallocAsync(&scratchH, (size_t)n * sizeof(double), s));
      hbuf = scratchH;
    }
    UB_CUDA(cudaMallocAsync(&scratchSt, (size_t)n * 2 * sizeof(uint32_t), s));
    uint32_t *i2gSt = scratchSt, *state = status ? status : scratchSt + n;
    cudaError_t e = launchFill(m->smCount, (long)n, h0, hbuf, state, s);
    for (int it = 0; it < max_iterations && e == cudaSuccess; it++) {
      e = launchImageToGround(*m, (long)n, line, samp, nullptr, hbuf, 1, prec, x, y, z, nullptr, i2gSt, s);
      if (e == cudaSuccess)
        e = launchDemHeightUpdate(d, m->smCount, (long)n, x, y, z, hbuf, i2gSt, state, tolerance,
                                  it == max_iterations - 1 ? 1 : 0, s);
    }
    if (scratchH) cudaFreeAsync(scratchH, s);
    cudaFreeAsync(scratchSt, s);
    UB_CUDA(e);
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_image_to_ground_dem(usgscsm_b200_model *m, const usgscsm_b200_dem *dem, int64_t n,
                                     const double *line, const double *samp, double h0,
                                     double tolerance, int max_iterations, double prec, double *x,
                                     double *y, double *z, double *height, uint32_t *status) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (int rc = checkDem(dem, 0, 0, dem ? dem->n_rows : 0, dem ? dem->n_cols : 0)) return rc;
  if (n == 0) return USGSCSM_B200_OK;
  if (!line || !samp || !x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  // the whole raster once, then the points through the pipeline
  void *devSamples = nullptr;
  const size_t bytes = (size_t)dem->n_rows * dem->n_cols * sampleBytes(dem->sample_type);
  int rc = guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    UB_CUDA(cudaMalloc(&devSamples, bytes ? bytes : 1));
    UB_CUDA(cudaMemcpy(devSamples, dem->samples, bytes, cudaMemcpyHostToDevice));
    return USGSCSM_B200_OK;
  });
  if (rc == USGSCSM_B200_OK) {
    usgscsm_b200_dem onDevice = *dem;
    onDevice.samples = devSamples;
    rc = runPipelined(m, n, {{line, 1, false}, {samp, 1, false}}, {{x, 1}, {y, 1}, {z, 1}, {height, 1}},
                      status, 0, [&](const Chunk &k) {
                        int r = usgscsm_b200_image_to_ground_dem_dev(
                            m, &onDevice, k.count, k.in[0], k.in[1], h0, tolerance, max_iterations, prec,
                            k.out[0], k.out[1], k.out[2], k.out[3], k.status, k.stream);
                        return r == USGSCSM_B200_OK ? cudaSuccess : cudaErrorUnknown;
                      });
  }
  if (devSamples) {
    cudaSetDevice(m->device);
    cudaFree(devSamples);
  }
  return rc;
}

}  // extern "C"
