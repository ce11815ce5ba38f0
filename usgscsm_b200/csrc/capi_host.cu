// capi_host.cu -- the HOST-pointer entry points of include/usgscsm_b200.h.
//
// Every one of them is the same machine: the arrays are cut into chunks, and each chunk goes
// through one of a few pipeline slots -- H2D copies, the kernel(s), D2H copies, all on the
// slot's own stream -- so that while chunk c computes, chunk c+1 is copied in and chunk c-1 is
// copied out.  Copies are truly asynchronous when the caller's arrays are page-locked
// (cudaHostAlloc / cudaHostRegister / torch pin_memory); with pageable arrays the runtime
// stages them and the calls still complete correctly.  The entry points differ only in which
// arrays they move and which launcher they call; that is what they state below.
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include "capi_common.h"
#include "pipeline.cuh"

namespace ub200 {

thread_local bool tlsSingleThreadPipeline = false;

bool pipelineThreads() {
  if (tlsSingleThreadPipeline) return false;
  static const bool on = [] {
    const char *v = getenv("USGSCSM_B200_PIPELINE_THREADS");
    return !(v && v[0] == '0');
  }();
  return on;
}
static long envLongHost(const char *name, long def, long lo, long hi) {
  const char *v = getenv(name);
  long x = v && *v ? atol(v) : def;
  return x < lo ? lo : (x > hi ? hi : x);
}
int stagedSlots() {
  static const int n = (int)envLongHost("USGSCSM_B200_STAGE_THREADS", 12, 1, kMaxPipelineSlots);
  return n;
}
long stagedChunk() {
  static const long n = envLongHost("USGSCSM_B200_STAGE_CHUNK", 1L << 18, 1024, 1L << 26);
  return n;
}

// True when p lives in ordinary (pageable) host memory, i.e. not in cudaHostAlloc /
// cudaHostRegister memory.
bool hostPointerPageable(const void *p) {
  if (!p) return false;
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return true;
  }
  return a.type == cudaMemoryTypeUnregistered;
}

}  // namespace ub200

using namespace ub200;

extern "C" {

int usgscsm_b200_ground_to_image(usgscsm_b200_model *m, int64_t n, const double *x,
                                 const double *y, const double *z, double prec, double *line,
                                 double *samp, double *achieved, uint32_t *status) {
  int rc;
  if (badCall(m, n, rc)) return rc;
  if (!x || !y || !z || !line || !samp) return fail(USGSCSM_B200_EINVAL, "null array");
  return runPipelined(m, n, {{x, 1, false}, {y, 1, false}, {z, 1, false}},
                      {{line, 1}, {samp, 1}, {achieved, 1}}, status, 0, [&](const Chunk &k) {
                        return launchGroundToImage(*m, k.count, k.in[0], k.in[1], k.in[2], prec,
                                                   k.out[0], k.out[1], k.out[2], k.status, k.stream);
                      });
}

int usgscsm_b200_image_to_ground(usgscsm_b200_model *m, int64_t n, const double *line,
                                 const double *samp, const double *height, int height_stride,
                                 double prec, double *x, double *y, double *z, double *achieved,
                                 uint32_t *status) {
  int rc;
  if (badCall(m, n, rc)) return rc;
  if (!line || !samp || !height || !x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  const int hs = height_stride ? 1 : 0;
  return runPipelined(m, n, {{line, 1, false}, {samp, 1, false}, {height, 1, hs == 0}},
                      {{x, 1}, {y, 1}, {z, 1}, {achieved, 1}}, status, 0, [&](const Chunk &k) {
                        return launchImageToGround(*m, k.count, k.in[0], k.in[1], nullptr, k.in[2],
                                                   hs, prec, k.out[0], k.out[1], k.out[2],
                                                   k.out[3], k.status, k.stream);
                      });
}

// Dense grid, host arrays: only the heights go up (8 B/pt, or nothing for a single height) and
// the ground points come down; the pixel coordinates are generated in the kernel.  Chunks are
// whole image rows.
int usgscsm_b200_image_to_ground_grid(usgscsm_b200_model *m, double line0, double samp0,
                                      double dline, double dsamp, int64_t n_lines,
                                      int64_t n_samps, const double *height, int height_stride,
                                      double prec, double *x, double *y, double *z,
                                      double *achieved, uint32_t *status) {
  if (!m || n_lines < 0 || n_samps < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n_lines == 0 || n_samps == 0) return USGSCSM_B200_OK;
  if (!height || !x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  const int hs = height_stride ? 1 : 0;
  return runPipelined(
      m, n_lines * n_samps, {{height, 1, hs == 0}}, {{x, 1}, {y, 1}, {z, 1}, {achieved, 1}}, status, 0,
      [&](const Chunk &k) {
        GridSpec g{line0, samp0, dline, dsamp, (long)n_samps, (long)(k.offset / n_samps)};
        return launchImageToGround(*m, k.count, nullptr, nullptr, &g, k.in[0], hs, prec, k.out[0],
                                   k.out[1], k.out[2], k.out[3], k.status, k.stream);
      },
      (long)n_samps);
}

// AoS flavours (arrays of csm::EcefCoord / csm::ImageCoord, csm/csm.h:212-308): a chunk is
// copied as it lies, split into SoA on the device, run through the same kernels, joined and
// copied back.
int usgscsm_b200_ground_to_image_aos(usgscsm_b200_model *m, int64_t n, const double *xyz,
                                     double prec, double *line_samp, double *achieved,
                                     uint32_t *status) {
  int rc;
  if (badCall(m, n, rc)) return rc;
  if (!xyz || !line_samp) return fail(USGSCSM_B200_EINVAL, "null array");
  return runPipelined(m, n, {{xyz, 3, false}}, {{line_samp, 2}, {achieved, 1}}, status, 5,
                      [&](const Chunk &k) {
                        double *const *w = k.work.data();  // x y z | line samp
                        cudaError_t e = launchAosSplit(3, k.count, k.in[0], w[0], w[1], w[2], k.stream);
                        if (e == cudaSuccess)
                          e = launchGroundToImage(*m, k.count, w[0], w[1], w[2], prec, w[3], w[4],
                                                  k.out[1], k.status, k.stream);
                        if (e == cudaSuccess)
                          e = launchAosJoin(2, k.count, w[3], w[4], nullptr, k.out[0], k.stream);
                        return e;
                      });
}

int usgscsm_b200_image_to_ground_aos(usgscsm_b200_model *m, int64_t n, const double *line_samp,
                                     const double *height, int height_stride, double prec,
                                     double *xyz, double *achieved, uint32_t *status) {
  int rc;
  if (badCall(m, n, rc)) return rc;
  if (!line_samp || !height || !xyz) return fail(USGSCSM_B200_EINVAL, "null array");
  const int hs = height_stride ? 1 : 0;
  return runPipelined(m, n, {{line_samp, 2, false}, {height, 1, hs == 0}}, {{xyz, 3}, {achieved, 1}},
                      status, 5, [&](const Chunk &k) {
                        double *const *w = k.work.data();  // line samp | x y z
                        cudaError_t e = launchAosSplit(2, k.count, k.in[0], w[0], w[1], nullptr, k.stream);
                        if (e == cudaSuccess)
                          e = launchImageToGround(*m, k.count, w[0], w[1], nullptr, k.in[1], hs, prec,
                                                  w[2], w[3], w[4], k.out[1], k.status, k.stream);
                        if (e == cudaSuccess)
                          e = launchAosJoin(3, k.count, w[2], w[3], w[4], k.out[0], k.stream);
                        return e;
                      });
}

int usgscsm_b200_image_to_remote_imaging_locus(usgscsm_b200_model *m, int64_t n, const double *line,
                                               const double *samp, double prec, double *px,
                                               double *py, double *pz, double *dx, double *dy,
                                               double *dz, double *achieved, uint32_t *status) {
  int rc;
  if (badCall(m, n, rc)) return rc;
  if (!line || !samp || !px || !py || !pz || !dx || !dy || !dz)
    return fail(USGSCSM_B200_EINVAL, "null array");
  return runPipelined(m, n, {{line, 1, false}, {samp, 1, false}},
                      {{px, 1}, {py, 1}, {pz, 1}, {dx, 1}, {dy, 1}, {dz, 1}, {achieved, 1}}, status,
                      0, [&](const Chunk &k) {
                        LocusOut o{k.out[0], k.out[1], k.out[2], k.out[3],
                                   k.out[4], k.out[5], k.out[6], k.status};
                        return launchRemoteLocus(*m, k.count, k.in[0], k.in[1], prec, o, k.stream);
                      });
}

int usgscsm_b200_image_to_proximate_imaging_locus(
    usgscsm_b200_model *m, int64_t n, const double *line, const double *samp, const double *x,
    const double *y, const double *z, double prec, double *px, double *py, double *pz, double *dx,
    double *dy, double *dz, double *achieved, uint32_t *status) {
  int rc;
  if (badCall(m, n, rc)) return rc;
  if (!line || !samp || !x || !y || !z || !px || !py || !pz || !dx || !dy || !dz)
    return fail(USGSCSM_B200_EINVAL, "null array");
  return runPipelined(
      m, n, {{line, 1, false}, {samp, 1, false}, {x, 1, false}, {y, 1, false}, {z, 1, false}},
      {{px, 1}, {py, 1}, {pz, 1}, {dx, 1}, {dy, 1}, {dz, 1}, {achieved, 1}}, status, 0,
      [&](const Chunk &k) {
        LocusOut o{k.out[0], k.out[1], k.out[2], k.out[3], k.out[4], k.out[5], k.out[6], k.status};
        return launchProximateLocus(*m, k.count, k.in[0], k.in[1], k.in[2], k.in[3], k.in[4], prec, o,
                                    k.stream);
      });
}

int usgscsm_b200_ground_partials(usgscsm_b200_model *m, int64_t n, const double *x, const double *y,
                                 const double *z, double *partials, uint32_t *status) {
  int rc;
  if (badCall(m, n, rc)) return rc;
  if (!x || !y || !z || !partials) return fail(USGSCSM_B200_EINVAL, "null array");
  return runPipelined(m, n, {{x, 1, false}, {y, 1, false}, {z, 1, false}}, {{partials, 6}}, status, 0,
                      [&](const Chunk &k) {
                        return launchGroundPartials(*m, k.count, k.in[0], k.in[1], k.in[2], k.out[0],
                                                    k.status, k.stream);
                      });
}

int usgscsm_b200_sensor_partials(usgscsm_b200_model *m, int index, int64_t n, const double *x,
                                 const double *y, const double *z, const double *img_line,
                                 const double *img_samp, double prec, double *dline, double *dsamp,
                                 double *achieved, uint32_t *status) {
  int c = sensorPartialsCheck(m, index, n, x, y, z, dline, dsamp);
  if (c != USGSCSM_B200_OK || n == 0) return c;
  const bool haveImage = img_line && img_samp;
  std::vector<HostIn> ins = {{x, 1, false}, {y, 1, false}, {z, 1, false}};
  if (haveImage) {
    ins.push_back({img_line, 1, false});
    ins.push_back({img_samp, 1, false});
  }
  // the combine step needs a device status array even when the caller wants none
  return runPipelined(m, n, ins, {{dline, 1}, {dsamp, 1}, {achieved, 1}}, status, 1,
                      [&](const Chunk &k) {
                        uint32_t *st = k.status ? k.status : reinterpret_cast<uint32_t *>(k.work[0]);
                        sensorPartialsDev(m, index, k.count, k.in[0], k.in[1], k.in[2],
                                          haveImage ? k.in[3] : nullptr, haveImage ? k.in[4] : nullptr,
                                          prec, k.out[0], k.out[1], k.out[2], st, k.stream);
                        return cudaSuccess;  // sensorPartialsDev throws CudaError on failure
                      });
}

int usgscsm_b200_all_sensor_partials(usgscsm_b200_model *m, int n_indices, const int *indices,
                                     int64_t n, const double *x, const double *y, const double *z,
                                     const double *img_line, const double *img_samp, double prec,
                                     double *partials, double *achieved, uint32_t *status) {
  int c = allSensorPartialsCheck(m, n_indices, indices, n, x, y, z, partials);
  if (c != USGSCSM_B200_OK || n == 0 || n_indices == 0) return c;
  const bool haveImage = img_line && img_samp;
  std::vector<HostIn> ins = {{x, 1, false}, {y, 1, false}, {z, 1, false}};
  if (haveImage) {
    ins.push_back({img_line, 1, false});
    ins.push_back({img_samp, 1, false});
  }
  return runPipelined(m, n, ins, {{partials, 2 * n_indices}, {achieved, 1}}, status, 0,
                      [&](const Chunk &k) {
                        allSensorPartialsDev(m, n_indices, indices, k.count, k.in[0], k.in[1],
                                             k.in[2], haveImage ? k.in[3] : nullptr,
                                             haveImage ? k.in[4] : nullptr, prec, k.out[0],
                                             k.out[1], k.status, k.stream);
                        return cudaSuccess;  // throws CudaError on failure
                      });
}

int usgscsm_b200_image_sensor_state(usgscsm_b200_model *m, int64_t n, const double *line,
                                    const double *samp, double *time, double *px, double *py,
                                    double *pz, double *vx, double *vy, double *vz,
                                    uint32_t *status) {
  int rc;
  if (badCall(m, n, rc)) return rc;
  if (!line || !samp) return fail(USGSCSM_B200_EINVAL, "null array");
  return runPipelined(m, n, {{line, 1, false}, {samp, 1, false}},
                      {{time, 1}, {px, 1}, {py, 1}, {pz, 1}, {vx, 1}, {vy, 1}, {vz, 1}}, status, 0,
                      [&](const Chunk &k) {
                        int r = usgscsm_b200_image_sensor_state_dev(
                            m, k.count, k.in[0], k.in[1], k.out[0], k.out[1], k.out[2], k.out[3],
                            k.out[4], k.out[5], k.out[6], k.status, k.stream);
                        return r == USGSCSM_B200_OK ? cudaSuccess : cudaErrorUnknown;
                      });
}

int usgscsm_b200_sensor_state_at_time(usgscsm_b200_model *m, int64_t n, const double *time,
                                      double *px, double *py, double *pz, double *vx, double *vy,
                                      double *vz, uint32_t *status) {
  int rc;
  if (badCall(m, n, rc)) return rc;
  if (!time) return fail(USGSCSM_B200_EINVAL, "null array");
  return runPipelined(m, n, {{time, 1, false}},
                      {{px, 1}, {py, 1}, {pz, 1}, {vx, 1}, {vy, 1}, {vz, 1}}, status, 0,
                      [&](const Chunk &k) {
                        int r = usgscsm_b200_sensor_state_at_time_dev(
                            m, k.count, k.in[0], k.out[0], k.out[1], k.out[2], k.out[3], k.out[4],
                            k.out[5], k.status, k.stream);
                        return r == USGSCSM_B200_OK ? cudaSuccess : cudaErrorUnknown;
                      });
}

int usgscsm_b200_illumination_direction(usgscsm_b200_model *m, int64_t n, const double *x,
                                        const double *y, const double *z, double *dx, double *dy,
                                        double *dz, uint32_t *status) {
  int rc;
  if (badCall(m, n, rc)) return rc;
  if (!x || !y || !z || !dx || !dy || !dz) return fail(USGSCSM_B200_EINVAL, "null array");
  return runPipelined(m, n, {{x, 1, false}, {y, 1, false}, {z, 1, false}},
                      {{dx, 1}, {dy, 1}, {dz, 1}}, status, 0, [&](const Chunk &k) {
                        illuminationDev(m, k.count, k.in[0], k.in[1], k.in[2], k.out[0], k.out[1],
                                        k.out[2], k.status, k.stream);
                        return cudaSuccess;  // throws on failure
                      });
}

// Error propagation (covar.cu): the covariances travel as arrays of 9 / 4 doubles per point.
int usgscsm_b200_ground_to_image_covar(usgscsm_b200_model *m, int64_t n, const double *x,
                                       const double *y, const double *z, const double *ground_cov,
                                       double prec, double *line, double *samp, double *image_cov,
                                       double *achieved, uint32_t *status) {
  int c = covarCheck(m, n);
  if (c != USGSCSM_B200_OK || n == 0) return c;
  if (!x || !y || !z || !ground_cov || !line || !samp || !image_cov)
    return fail(USGSCSM_B200_EINVAL, "null array");
  return runPipelined(m, n, {{x, 1, false}, {y, 1, false}, {z, 1, false}, {ground_cov, 9, false}},
                      {{line, 1}, {samp, 1}, {image_cov, 4}, {achieved, 1}}, status, 1,
                      [&](const Chunk &k) {
                        uint32_t *st = k.status ? k.status : reinterpret_cast<uint32_t *>(k.work[0]);
                        groundToImageCovarDev(m, k.count, k.in[0], k.in[1], k.in[2], k.in[3], prec,
                                              k.out[0], k.out[1], k.out[2], k.out[3], st, k.stream);
                        return cudaSuccess;  // throws CudaError on failure
                      });
}

int usgscsm_b200_image_to_ground_covar(usgscsm_b200_model *m, int64_t n, const double *line,
                                       const double *samp, const double *image_cov,
                                       const double *height, int height_stride,
                                       const double *height_variance, int variance_stride,
                                       double prec, double *x, double *y, double *z,
                                       double *ground_cov, double *achieved, uint32_t *status) {
  int c = covarCheck(m, n);
  if (c != USGSCSM_B200_OK || n == 0) return c;
  if (!line || !samp || !image_cov || !height || !height_variance || !x || !y || !z || !ground_cov)
    return fail(USGSCSM_B200_EINVAL, "null array");
  const int hs = height_stride ? 1 : 0, vs = variance_stride ? 1 : 0;
  return runPipelined(
      m, n,
      {{line, 1, false}, {samp, 1, false}, {image_cov, 4, false}, {height, 1, hs == 0},
       {height_variance, 1, vs == 0}},
      {{x, 1}, {y, 1}, {z, 1}, {ground_cov, 9}, {achieved, 1}}, status, 1, [&](const Chunk &k) {
        uint32_t *st = k.status ? k.status : reinterpret_cast<uint32_t *>(k.work[0]);
        imageToGroundCovarDev(m, k.count, k.in[0], k.in[1], k.in[2], k.in[3], hs, k.in[4], vs, prec,
                              k.out[0], k.out[1], k.out[2], k.out[3], k.out[4], st, k.stream);
        return cudaSuccess;
      });
}

}  // extern "C"

// ---- several GPUs from one process -----------------------------------------------------------
// Points are independent and the packed state is tiny, so a batch shards with no exchange
// step: contiguous chunk g of the arrays goes to models[g] (the same state
// loaded on different devices; sizes differ by at most one point), one host thread per model,
// each running the pipeline above on its own device.
namespace {

void splitChunk(int64_t n, int parts, int g, int64_t &off, int64_t &cnt) {
  const int64_t base = n / parts, extra = n % parts;
  off = base * g + (g < extra ? g : extra);
  cnt = base + (g < extra ? 1 : 0);
}

template <class F>
int runOnModels(usgscsm_b200_model *const *models, int n_models, int64_t n, F &&call) {
  if (!models || n_models < 1 || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  for (int g = 0; g < n_models; g++) {
    if (!models[g]) return fail(USGSCSM_B200_EINVAL, "null model");
    for (int k = 0; k < g; k++)
      if (models[k] == models[g]) return fail(USGSCSM_B200_EINVAL, "a model is listed twice");
  }
  std::vector<int> codes((size_t)n_models, USGSCSM_B200_OK);
  std::vector<std::string> messages((size_t)n_models);
  std::vector<std::thread> workers;
  for (int g = 0; g < n_models; g++) {
    workers.emplace_back([&, g]() {
      int64_t off, cnt;
      splitChunk(n, n_models, g, off, cnt);
      tlsSingleThreadPipeline = true;
      codes[g] = call(models[g], off, cnt);
      if (codes[g] != USGSCSM_B200_OK) messages[g] = lastError();  // the worker thread's own copy
    });
  }
  for (std::thread &t : workers) t.join();
  for (int g = 0; g < n_models; g++)
    if (codes[g] != USGSCSM_B200_OK) {
      lastError() = "model " + std::to_string(g) + ": " + messages[g];
      return codes[g];
    }
  return USGSCSM_B200_OK;
}

}  // namespace

extern "C" {

int usgscsm_b200_ground_to_image_multi(usgscsm_b200_model *const *models, int n_models, int64_t n,
                                       const double *x, const double *y, const double *z,
                                       double prec, double *line, double *samp, double *achieved,
                                       uint32_t *status) {
  if (n > 0 && (!x || !y || !z || !line || !samp)) return fail(USGSCSM_B200_EINVAL, "null array");
  return runOnModels(models, n_models, n, [&](usgscsm_b200_model *m, int64_t off, int64_t cnt) {
    return usgscsm_b200_ground_to_image(m, cnt, x + off, y + off, z + off, prec, line + off,
                                        samp + off, achieved ? achieved + off : nullptr,
                                        status ? status + off : nullptr);
  });
}

int usgscsm_b200_image_to_ground_multi(usgscsm_b200_model *const *models, int n_models, int64_t n,
                                       const double *line, const double *samp,
                                       const double *height, int height_stride, double prec,
                                       double *x, double *y, double *z, double *achieved,
                                       uint32_t *status) {
  if (n > 0 && (!line || !samp || !height || !x || !y || !z))
    return fail(USGSCSM_B200_EINVAL, "null array");
  return runOnModels(models, n_models, n, [&](usgscsm_b200_model *m, int64_t off, int64_t cnt) {
    return usgscsm_b200_image_to_ground(m, cnt, line + off, samp + off,
                                        height_stride ? height + off : height, height_stride,
                                        prec, x + off, y + off, z + off,
                                        achieved ? achieved + off : nullptr,
                                        status ? status + off : nullptr);
  });
}

}  // extern "C"
