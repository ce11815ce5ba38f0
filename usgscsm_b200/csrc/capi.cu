// capi.cu -- the C-ABI of include/usgscsm_b200.h over the kernel launchers.
//
// There is no CPU implementation of the per-point math behind these entry points: every
// compute call ends in a kernel launch or fails with a negative error code.
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/usgscsm_b200.h"
#include "capi_common.h"
#include "model.h"
#include "state_reader.h"
#include "work_queue.cuh"

namespace ub200 {

static std::atomic<int64_t> g_launches{0};
void countLaunch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static thread_local std::string g_lastError;

static long envLong(const char *name, long def, long lo, long hi) {
  const char *v = getenv(name);
  long x = v && *v ? atol(v) : def;
  return x < lo ? lo : (x > hi ? hi : x);
}
int pipelineSlots() {
  static const int n = (int)envLong("USGSCSM_B200_PIPELINE_SLOTS", kDefaultPipelineSlots, 1, kMaxPipelineSlots);
  return n;
}
long pipelineChunk() {
  static const long n = envLong("USGSCSM_B200_PIPELINE_CHUNK", kDefaultPipelineChunk, 1024, 1L << 28);
  return n;
}

int grabSize() {
  static const int grab = [] {
    const char *v = getenv("USGSCSM_B200_GRAB");
    int g = v && *v ? atoi(v) : kDefaultGrab;
    if (g < 32) g = 32;
    return (g + 31) / 32 * 32;
  }();
  return grab;
}

// ---- Model ------------------------------------------------------------------------------
Model::~Model() {
  cudaSetDevice(device);
  for (DeviceBuffer &b : owned) cudaFree(b.ptr);
  for (GeoTables &t : geoTables) cudaFree(t.dev);
  for (PipelineSlot &s : slots) {
    cudaFree(s.scratch);
    if (s.stage) cudaFreeHost(s.stage);
    if (s.done) cudaEventDestroy(s.done);
    if (s.stream) cudaStreamDestroy(s.stream);
  }
}

const double *Model::upload(const std::vector<double> &v) {
  DeviceBuffer b;
  b.bytes = (v.empty() ? 1 : v.size()) * sizeof(double);
  UB_CUDA(cudaMalloc(&b.ptr, b.bytes));
  owned.push_back(b);
  if (!v.empty()) UB_CUDA(cudaMemcpy(b.ptr, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice));
  return static_cast<const double *>(b.ptr);
}

void Model::ensureScratch(int s, long doubles) {
  PipelineSlot &p = slots[s];
  if (!p.stream) UB_CUDA(cudaStreamCreateWithFlags(&p.stream, cudaStreamNonBlocking));
  if (!p.done) UB_CUDA(cudaEventCreateWithFlags(&p.done, cudaEventDisableTiming));
  if (p.scratchDoubles >= doubles) return;
  cudaFree(p.scratch);
  p.scratch = nullptr;
  p.scratchDoubles = 0;
  UB_CUDA(cudaMalloc(&p.scratch, doubles * sizeof(double)));
  p.scratchDoubles = doubles;
}

void Model::ensureStage(int s, long bytes) {
  PipelineSlot &p = slots[s];
  if (p.stageBytes >= bytes) return;
  if (p.stage) cudaFreeHost(p.stage);
  p.stage = nullptr;
  p.stageBytes = 0;
  UB_CUDA(cudaHostAlloc(&p.stage, (size_t)bytes, cudaHostAllocDefault));
  p.stageBytes = bytes;
}

// ---- error plumbing (capi_common.h) --------------------------------------------------------
std::string &lastError() { return g_lastError; }

int deviceCount() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

// Dispatch of the two per-point operations on the model kind (device pointers).
cudaError_t launchGroundToImage(Model &m, long n, const double *x, const double *y,
                                       const double *z, double prec, double *line, double *samp,
                                       double *achieved, uint32_t *status, cudaStream_t stream) {
  switch (m.kind) {
    case USGSCSM_B200_MODEL_LINE_SCANNER:
      return launchLsGroundToImage(m.ls, m.smCount, n, x, y, z, prec, line, samp, achieved,
                                   status, stream);
    case USGSCSM_B200_MODEL_PUSH_FRAME:
      return launchPfGroundToImage(m.ls, m.smCount, n, x, y, z, prec, line, samp, achieved,
                                   status, stream);
    case USGSCSM_B200_MODEL_FRAME:
      return launchFrGroundToImage(m.fr, m.smCount, n, x, y, z, prec, line, samp, achieved,
                                   status, stream);
    case USGSCSM_B200_MODEL_SAR:
      return launchSarGroundToImage(m.sar, m.smCount, n, x, y, z, prec, line, samp, achieved,
                                    status, stream);
    default:
      return cudaErrorNotSupported;
  }
}

cudaError_t launchImageToGround(Model &m, long n, const double *line, const double *samp,
                                       const GridSpec *grid, const double *height,
                                       int heightStride, double prec, double *x, double *y,
                                       double *z, double *achieved, uint32_t *status,
                                       cudaStream_t stream) {
  switch (m.kind) {
    case USGSCSM_B200_MODEL_LINE_SCANNER:
      return launchLsImageToGround(m.ls, m.smCount, n, line, samp, grid, height, heightStride,
                                   prec, x, y, z, achieved, status, stream);
    case USGSCSM_B200_MODEL_PUSH_FRAME:
      return launchPfImageToGround(m.ls, m.smCount, n, line, samp, grid, height, heightStride,
                                   prec, x, y, z, achieved, status, stream);
    case USGSCSM_B200_MODEL_FRAME:
      return launchFrImageToGround(m.fr, m.smCount, n, line, samp, grid, height, heightStride,
                                   prec, x, y, z, achieved, status, stream);
    case USGSCSM_B200_MODEL_SAR:
      return launchSarImageToGround(m.sar, m.smCount, n, line, samp, grid, height, heightStride,
                                    prec, x, y, z, achieved, status, stream);
    default:
      return cudaErrorNotSupported;
  }
}

}  // namespace ub200

namespace ub200 {
void checkDistortionCoeffs(int type, size_t n);  // model_ls.cu
namespace isd {  // isd_to_state.cpp
std::string stateFromIsd(const char *isdText, size_t len, const std::string &modelName,
                         const std::string &imageIdentifier);
}
}
using namespace ub200;

extern "C" {

int usgscsm_b200_abi_version(void) { return USGSCSM_B200_ABI_VERSION; }
const char *usgscsm_b200_last_error(void) { return g_lastError.c_str(); }
int usgscsm_b200_device_count(void) { return deviceCount(); }
int64_t usgscsm_b200_kernel_launches(void) { return g_launches.load(); }

// getModelNameFromModelState (src/UsgsAstroPlugin.cpp:176-218): the JSON key "m_modelName".
int usgscsm_b200_model_from_state(const char *state, size_t len, int device,
                                  usgscsm_b200_model **out) {
  if (!state || !out) return fail(USGSCSM_B200_EINVAL, "null argument");
  *out = nullptr;
  if (device < 0 || device >= deviceCount())
    return fail(USGSCSM_B200_ENODEVICE, "no usable CUDA device with that index");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(device));
    json::Value root = parseState(state, len);
    StateReader st(root);
    std::unique_ptr<usgscsm_b200_model> m(new usgscsm_b200_model);
    m->device = device;
    cudaDeviceProp prop;
    UB_CUDA(cudaGetDeviceProperties(&prop, device));
    m->smCount = prop.multiProcessorCount;
    {
      // Stream-ordered scratch (dense-grid row / column tables) comes from the device's default
      // memory pool; keep freed blocks cached instead of returning them to the driver at every
      // synchronisation (the default threshold of 0 makes each cudaMallocAsync a driver call).
      cudaMemPool_t pool;
      UB_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
      uint64_t threshold = ~0ull;
      UB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold));
    }
    m->modelName = st.getString("m_modelName");
    // before the loaders: updateState replaces it for states that ask to be derived (LS:283-285)
    m->covariance = st.getVecDouble("m_covariance", std::vector<double>());
    if (m->modelName == "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL") {
      m->kind = USGSCSM_B200_MODEL_LINE_SCANNER;
      loadLineScanner(*m, st);
    } else if (m->modelName == "USGS_ASTRO_PUSH_FRAME_SENSOR_MODEL") {
      m->kind = USGSCSM_B200_MODEL_PUSH_FRAME;
      loadPushFrame(*m, st);
    } else if (m->modelName == "USGS_ASTRO_FRAME_SENSOR_MODEL") {
      m->kind = USGSCSM_B200_MODEL_FRAME;
      loadFrame(*m, st);
    } else if (m->modelName == "USGS_ASTRO_SAR_SENSOR_MODEL") {
      m->kind = USGSCSM_B200_MODEL_SAR;
      loadSar(*m, st);
    } else {
      g_lastError = "unsupported model name: " + m->modelName;
      return USGSCSM_B200_EUNSUPPORTED;
    }
    loadSun(*m, st);
    m->stateTree = root;
    UB_CUDA(cudaDeviceSynchronize());
    *out = m.release();
    return USGSCSM_B200_OK;
  });
}

// constructStateFromIsd of the four models (isd_to_state.cpp).  Host only.
int usgscsm_b200_state_from_isd(const char *isd, size_t len, const char *model_name,
                                const char *image_id, char *out, size_t capacity,
                                size_t *needed) {
  if (!isd || !needed) return fail(USGSCSM_B200_EINVAL, "null argument");
  return guarded([&]() -> int {
    const std::string text =
        isd::stateFromIsd(isd, len, model_name ? model_name : "", image_id ? image_id : "");
    *needed = text.size();
    if (out && capacity >= text.size()) std::memcpy(out, text.data(), text.size());
    else if (out) return fail(USGSCSM_B200_EINVAL, "output buffer too small");
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_model_from_isd(const char *isd, size_t len, const char *model_name,
                                const char *image_id, int device, usgscsm_b200_model **out) {
  if (!isd || !out) return fail(USGSCSM_B200_EINVAL, "null argument");
  *out = nullptr;
  std::string text;
  int rc = guarded([&]() -> int {
    text = isd::stateFromIsd(isd, len, model_name ? model_name : "", image_id ? image_id : "");
    return USGSCSM_B200_OK;
  });
  if (rc != USGSCSM_B200_OK) return rc;
  return usgscsm_b200_model_from_state(text.data(), text.size(), device, out);
}

// getModelNameFromModelState (src/UsgsAstroPlugin.cpp:176-218).  Host only: no device needed.
int usgscsm_b200_model_name_from_state(const char *state, size_t len, char *name, size_t cap) {
  if (!state || !name || cap == 0) return fail(USGSCSM_B200_EINVAL, "null argument");
  name[0] = 0;
  return guarded([&]() -> int {
    json::Value root = parseState(state, len);
    StateReader st(root);
    std::string n = st.getString("m_modelName");
    if (n != "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL" && n != "USGS_ASTRO_PUSH_FRAME_SENSOR_MODEL" &&
        n != "USGS_ASTRO_FRAME_SENSOR_MODEL" && n != "USGS_ASTRO_SAR_SENSOR_MODEL") {
      g_lastError = "unsupported model name: " + n;
      return USGSCSM_B200_EUNSUPPORTED;
    }
    if (n.size() + 1 > cap) return fail(USGSCSM_B200_EINVAL, "name buffer too small");
    std::memcpy(name, n.c_str(), n.size() + 1);
    return USGSCSM_B200_OK;
  });
}

// nlohmann::json::to_msgpack(stateAsJson(state)): bin/usgscsm_cam_test.cc
// --output-binary-model-state.  Host only.
int usgscsm_b200_state_to_msgpack(const char *state, size_t len, unsigned char *out, size_t cap,
                                  size_t *needed) {
  if (!state || !needed) return fail(USGSCSM_B200_EINVAL, "null argument");
  return guarded([&]() -> int {
    std::string bytes = msgpack::dump(parseState(state, len));
    *needed = bytes.size();
    if (out && cap >= bytes.size()) std::memcpy(out, bytes.data(), bytes.size());
    else if (out) return fail(USGSCSM_B200_EINVAL, "output buffer too small");
    return USGSCSM_B200_OK;
  });
}

// The other direction: any accepted encoding (msgpack, JSON text, .sup) -> "<model name>\n<JSON>",
// the text form getModelState writes (what json::from_msgpack + dump give in the reference,
// bin/usgscsm_cam_test.cc:157-170).  No device needed.
int usgscsm_b200_state_to_json(const char *state, size_t len, char *out, size_t cap,
                               size_t *needed) {
  if (!state || !needed) return fail(USGSCSM_B200_EINVAL, "null argument");
  return guarded([&]() -> int {
    json::Value root = parseState(state, len);
    StateReader st(root);
    const std::string text = st.getString("m_modelName") + "\n" + json::dump(root);
    *needed = text.size();
    if (out && cap >= text.size()) std::memcpy(out, text.data(), text.size());
    else if (out) return fail(USGSCSM_B200_EINVAL, "output buffer too small");
    return USGSCSM_B200_OK;
  });
}

void usgscsm_b200_model_free(usgscsm_b200_model *m) { delete m; }

int usgscsm_b200_model_kind(const usgscsm_b200_model *m) { return m ? m->kind : 0; }
int usgscsm_b200_model_device(const usgscsm_b200_model *m) { return m ? m->device : -1; }

int usgscsm_b200_model_image_size(const usgscsm_b200_model *m, int *lines, int *samps) {
  if (!m || !lines || !samps) return fail(USGSCSM_B200_EINVAL, "null argument");
  switch (m->kind) {
    case USGSCSM_B200_MODEL_LINE_SCANNER:
      *lines = m->ls.nLines;
      *samps = m->ls.nSamples;
      break;
    case USGSCSM_B200_MODEL_PUSH_FRAME:  // getImageSize PF:1447-1449
      *lines = m->ls.nReducedLines;
      *samps = m->ls.nSamples;
      break;
    case USGSCSM_B200_MODEL_FRAME:
      *lines = m->fr.nLines;
      *samps = m->fr.nSamples;
      break;
    default:
      *lines = m->sar.nLines;
      *samps = m->sar.nSamples;
      break;
  }
  return USGSCSM_B200_OK;
}

int usgscsm_b200_model_num_parameters(const usgscsm_b200_model *m) {
  return m ? (int)m->params.size() : 0;
}

int usgscsm_b200_model_get_parameter(const usgscsm_b200_model *m, int index, double *value) {
  if (!m || !value) return fail(USGSCSM_B200_EINVAL, "null argument");
  if (index < 0 || index >= (int)m->params.size())
    return fail(USGSCSM_B200_EINVAL, "parameter index out of range");
  *value = m->params[index];
  return USGSCSM_B200_OK;
}

// getModelState (LS:345-483, PF:346-455, FR:1241-1300, SAR:395-455): the state the model was
// built from with the members a caller can change written back -- parameter values, parameter
// covariance, ellipsoid radii -- as "<model name>\n<JSON>".
int usgscsm_b200_model_get_state(const usgscsm_b200_model *m, char *out, size_t capacity,
                                 size_t *needed) {
  if (!m || !needed) return fail(USGSCSM_B200_EINVAL, "null argument");
  return guarded([&]() -> int {
    json::Value tree = m->stateTree;
    // unchanged entries keep their literal kind (integer / real), so an untouched model
    // re-serialises to the document it was built from
    auto putArray = [&](const char *key, const std::vector<double> &nums) {
      json::Value &dst = tree.members[key];
      if (dst.type == json::Type::Array && dst.numericArray && dst.nums.size() == nums.size()) {
        for (size_t k = 0; k < nums.size(); k++)
          if (std::memcmp(&dst.nums[k], &nums[k], sizeof(double)) != 0) {
            dst.nums[k] = nums[k];
            dst.numIsInt[k] = 0;
          }
      } else {
        dst = json::numberArray(nums);
      }
    };
    auto putNumber = [&](const char *key, double v) {
      json::Value &dst = tree.members[key];
      if (!(dst.isNumber() && std::memcmp(&dst.num, &v, sizeof(double)) == 0)) dst = json::number(v);
    };
    putArray("m_currentParameterValue", m->params);
    if (!m->covariance.empty()) putArray("m_covariance", m->covariance);
    const Ellipsoid &e = m->kind == USGSCSM_B200_MODEL_FRAME ? m->fr.ell
                         : m->kind == USGSCSM_B200_MODEL_SAR ? m->sar.ell : m->ls.ell;
    putNumber("m_majorAxis", e.majorAxis);
    putNumber("m_minorAxis", e.minorAxis);
    if (m->kind == USGSCSM_B200_MODEL_LINE_SCANNER || m->kind == USGSCSM_B200_MODEL_PUSH_FRAME) {
      // what updateState derived at load (LS:649-735), so that a reload does not derive it
      // again from the adjusted model
      putNumber("m_gsd", m->ls.gsd);
      putNumber("m_flyingHeight", m->ls.flyingHeight);
      putNumber("m_halfSwath", m->ls.halfSwath);
      putNumber("m_halfTime", m->ls.halfTime);
      putArray("m_referencePointXyz", std::vector<double>(m->referencePoint, m->referencePoint + 3));
    }
    const std::string text = m->modelName + "\n" + json::dump(tree);
    *needed = text.size();
    if (out && capacity >= text.size()) std::memcpy(out, text.data(), text.size());
    else if (out) return fail(USGSCSM_B200_EINVAL, "buffer too small");
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_model_set_parameter(usgscsm_b200_model *m, int index, double value) {
  if (!m) return fail(USGSCSM_B200_EINVAL, "null argument");
  if (index < 0 || index >= (int)m->params.size())
    return fail(USGSCSM_B200_EINVAL, "parameter index out of range");
  m->params[index] = value;
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    if (m->kind == USGSCSM_B200_MODEL_LINE_SCANNER) repackLineScanner(*m);
    if (m->kind == USGSCSM_B200_MODEL_PUSH_FRAME) repackPushFrame(*m);
    if (m->kind == USGSCSM_B200_MODEL_FRAME) repackFrame(*m);
    if (m->kind == USGSCSM_B200_MODEL_SAR) repackSar(*m);
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_model_ls_derived(const usgscsm_b200_model *m, double *out) {
  if (!m || !out) return fail(USGSCSM_B200_EINVAL, "null argument");
  if (m->kind != USGSCSM_B200_MODEL_LINE_SCANNER)
    return fail(USGSCSM_B200_EINVAL, "not a line-scan model");
  out[0] = m->ls.gsd;
  out[1] = m->ls.flyingHeight;
  out[2] = m->ls.halfSwath;
  out[3] = m->ls.halfTime;
  out[4] = m->ls.useApproxInitTrans ? 1.0 : 0.0;
  for (int i = 0; i < 14; i++) out[5 + i] = m->ls.projCoeffs[i];
  for (int i = 0; i < 3; i++) out[19 + i] = m->referencePoint[i];
  return USGSCSM_B200_OK;
}

int usgscsm_b200_model_ls_set_projective(usgscsm_b200_model *m, int use, const double *coeffs14) {
  if (!m) return fail(USGSCSM_B200_EINVAL, "null argument");
  if (m->kind != USGSCSM_B200_MODEL_LINE_SCANNER)
    return fail(USGSCSM_B200_EINVAL, "not a line-scan model");
  m->ls.useApproxInitTrans = use != 0;
  if (coeffs14) std::memcpy(m->ls.projCoeffs, coeffs14, 14 * sizeof(double));
  return USGSCSM_B200_OK;
}

// ---- device-pointer flavours ----------------------------------------------------------------
int usgscsm_b200_ground_to_image_dev(usgscsm_b200_model *m, int64_t n, const double *x,
                                     const double *y, const double *z, double prec, double *line,
                                     double *samp, double *achieved, uint32_t *status,
                                     void *stream) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!x || !y || !z || !line || !samp) return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    UB_CUDA(launchGroundToImage(*m, (long)n, x, y, z, prec, line, samp, achieved, status,
                                static_cast<cudaStream_t>(stream)));
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_image_to_ground_dev(usgscsm_b200_model *m, int64_t n, const double *line,
                                     const double *samp, const double *height, int height_stride,
                                     double prec, double *x, double *y, double *z,
                                     double *achieved, uint32_t *status, void *stream) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!line || !samp || !height || !x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    UB_CUDA(launchImageToGround(*m, (long)n, line, samp, nullptr, height, height_stride ? 1 : 0,
                                prec, x, y, z, achieved, status,
                                static_cast<cudaStream_t>(stream)));
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_image_to_ground_grid_dev(usgscsm_b200_model *m, double line0, double samp0,
                                          double dline, double dsamp, int64_t n_lines,
                                          int64_t n_samps, const double *height,
                                          int height_stride, double prec, double *x, double *y,
                                          double *z, double *achieved, uint32_t *status,
                                          void *stream) {
  if (!m || n_lines < 0 || n_samps < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n_lines == 0 || n_samps == 0) return USGSCSM_B200_OK;
  if (!height || !x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    GridSpec g{line0, samp0, dline, dsamp, (long)n_samps, 0};
    UB_CUDA(launchImageToGround(*m, (long)(n_lines * n_samps), nullptr, nullptr, &g, height,
                                height_stride ? 1 : 0, prec, x, y, z, achieved, status,
                                static_cast<cudaStream_t>(stream)));
    return USGSCSM_B200_OK;
  });
}

// The dense-grid imageToGround of the line-scan model without any allocation: the row / column
// tables go to a file-scope device array (ls_kernels.cu) instead of stream-ordered scratch.
int usgscsm_b200_image_to_ground_grid_static_dev(usgscsm_b200_model *m, double line0, double samp0,
                                                 double dline, double dsamp, int64_t n_lines,
                                                 int64_t n_samps, const double *height,
                                                 int height_stride, double prec, double *x, double *y,
                                                 double *z, double *achieved, uint32_t *status,
                                                 void *stream) {
  if (!m || n_lines < 0 || n_samps < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (m->kind != USGSCSM_B200_MODEL_LINE_SCANNER)
    return fail(USGSCSM_B200_EUNSUPPORTED,
                "usgscsm_b200_image_to_ground_grid_static_dev: line-scan models only (the other "
                "models' dense-grid kernels use no scratch: usgscsm_b200_image_to_ground_grid_dev "
                "allocates nothing for them)");
  if (n_lines == 0 || n_samps == 0) return USGSCSM_B200_OK;
  if (!height || !x || !y || !z) return fail(USGSCSM_B200_EINVAL, "null array");
  if (!staticGridScratchFits((long)n_lines, (long)n_samps))
    return fail(USGSCSM_B200_EINVAL,
                "usgscsm_b200_image_to_ground_grid_static_dev: grid too large for the static scratch "
                "(n_lines * 21 + n_samps * 4 must not exceed 2097152 doubles = 16 MiB); use "
                "usgscsm_b200_image_to_ground_grid_dev or split the grid into row bands");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    GridSpec g{line0, samp0, dline, dsamp, (long)n_samps, 0};
    UB_CUDA(launchLsImageToGroundGridStatic(m->ls, m->smCount, g, (long)n_lines, height,
                                            height_stride ? 1 : 0, prec, x, y, z, achieved, status,
                                            static_cast<cudaStream_t>(stream)));
    return USGSCSM_B200_OK;
  });
}

}  // extern "C"

// ---- many small images per call ------------------------------------------------------------------
namespace ub200 {
namespace {

// A few side streams per device, forked from and joined to the caller's stream: the kernels of
// consecutive images then overlap (one's tail with the next one's body).
constexpr int kSideStreams = 4;
struct SideStreams {
  cudaStream_t s[kSideStreams] = {};
  cudaEvent_t done[kSideStreams] = {};
  cudaEvent_t fork = nullptr;
  std::mutex mu;  // one fork / join at a time per device
};

SideStreams &sideStreams(int device) {
  static std::mutex mu;
  // never destroyed (CUDA objects must not be released from a static destructor)
  static std::vector<std::unique_ptr<SideStreams>> &perDevice = *new std::vector<std::unique_ptr<SideStreams>>();
  std::lock_guard<std::mutex> lock(mu);
  if ((int)perDevice.size() <= device) perDevice.resize((size_t)device + 1);
  if (!perDevice[(size_t)device]) {
    std::unique_ptr<SideStreams> p(new SideStreams);
    UB_CUDA(cudaEventCreateWithFlags(&p->fork, cudaEventDisableTiming));
    for (int k = 0; k < kSideStreams; k++) {
      UB_CUDA(cudaStreamCreateWithFlags(&p->s[k], cudaStreamNonBlocking));
      UB_CUDA(cudaEventCreateWithFlags(&p->done[k], cudaEventDisableTiming));
    }
    perDevice[(size_t)device] = std::move(p);
  }
  return *perDevice[(size_t)device];
}

// launchOne(k, offset, side stream) enqueues image k.
template <class F>
int forEachImage(usgscsm_b200_model *const *models, int n_models, const int64_t *counts,
                 cudaStream_t stream, F &&launchOne) {
  if (!models || n_models < 0 || (n_models > 0 && !counts)) return fail(USGSCSM_B200_EINVAL, "bad argument");
  for (int k = 0; k < n_models; k++) {
    if (!models[k] || counts[k] < 0) return fail(USGSCSM_B200_EINVAL, "null model / negative count");
    if (models[k]->device != models[0]->device)
      return fail(USGSCSM_B200_EINVAL, "the images of one call must live on one device");
  }
  if (n_models == 0) return USGSCSM_B200_OK;
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(models[0]->device));
    SideStreams &ss = sideStreams(models[0]->device);
    std::lock_guard<std::mutex> lock(ss.mu);
    const int used = n_models < kSideStreams ? n_models : kSideStreams;
    UB_CUDA(cudaEventRecord(ss.fork, stream));
    for (int j = 0; j < used; j++) UB_CUDA(cudaStreamWaitEvent(ss.s[j], ss.fork, 0));
    int64_t offset = 0;
    cudaError_t e = cudaSuccess;
    for (int k = 0; k < n_models && e == cudaSuccess; k++) {
      if (counts[k] > 0) e = launchOne(k, offset, ss.s[k % used]);
      offset += counts[k];
    }
    for (int j = 0; j < used; j++) {
      UB_CUDA(cudaEventRecord(ss.done[j], ss.s[j]));
      UB_CUDA(cudaStreamWaitEvent(stream, ss.done[j], 0));
    }
    UB_CUDA(e);
    return USGSCSM_B200_OK;
  });
}

}  // namespace
}  // namespace ub200

extern "C" {

int usgscsm_b200_ground_to_image_images_dev(usgscsm_b200_model *const *models, int n_models,
                                            const int64_t *counts, const double *x, const double *y,
                                            const double *z, double prec, double *line, double *samp,
                                            double *achieved, uint32_t *status, void *stream) {
  if (n_models > 0 && (!x || !y || !z || !line || !samp)) return fail(USGSCSM_B200_EINVAL, "null array");
  return forEachImage(models, n_models, counts, static_cast<cudaStream_t>(stream),
                      [&](int k, int64_t o, cudaStream_t s) {
                        return launchGroundToImage(*models[k], (long)counts[k], x + o, y + o, z + o, prec,
                                                   line + o, samp + o, achieved ? achieved + o : nullptr,
                                                   status ? status + o : nullptr, s);
                      });
}

int usgscsm_b200_image_to_ground_images_dev(usgscsm_b200_model *const *models, int n_models,
                                            const int64_t *counts, const double *line,
                                            const double *samp, const double *height,
                                            int height_stride, double prec, double *x, double *y,
                                            double *z, double *achieved, uint32_t *status, void *stream) {
  if (n_models > 0 && (!line || !samp || !height || !x || !y || !z))
    return fail(USGSCSM_B200_EINVAL, "null array");
  const int hs = height_stride ? 1 : 0;
  return forEachImage(models, n_models, counts, static_cast<cudaStream_t>(stream),
                      [&](int k, int64_t o, cudaStream_t s) {
                        return launchImageToGround(*models[k], (long)counts[k], line + o, samp + o, nullptr,
                                                   hs ? height + o : height, hs, prec, x + o, y + o, z + o,
                                                   achieved ? achieved + o : nullptr,
                                                   status ? status + o : nullptr, s);
                      });
}

int usgscsm_b200_image_to_ground_grid_images_dev(usgscsm_b200_model *const *models, int n_models,
                                                 double line0, double samp0, double dline, double dsamp,
                                                 const int64_t *n_lines, const int64_t *n_samps,
                                                 const double *height, int height_stride, double prec,
                                                 double *x, double *y, double *z, double *achieved,
                                                 uint32_t *status, void *stream) {
  if (n_models < 0 || (n_models > 0 && (!n_lines || !n_samps || !height || !x || !y || !z)))
    return fail(USGSCSM_B200_EINVAL, "bad argument");
  std::vector<int64_t> counts((size_t)(n_models > 0 ? n_models : 0));
  for (int k = 0; k < n_models; k++) {
    if (n_lines[k] < 0 || n_samps[k] < 0) return fail(USGSCSM_B200_EINVAL, "negative grid size");
    counts[(size_t)k] = n_lines[k] * n_samps[k];
  }
  const int hs = height_stride ? 1 : 0;
  return forEachImage(models, n_models, counts.data(), static_cast<cudaStream_t>(stream),
                      [&](int k, int64_t o, cudaStream_t s) {
                        GridSpec g{line0, samp0, dline, dsamp, (long)n_samps[k], 0};
                        return launchImageToGround(*models[k], (long)counts[(size_t)k], nullptr, nullptr, &g,
                                                   hs ? height + o : height, hs, prec, x + o, y + o, z + o,
                                                   achieved ? achieved + o : nullptr,
                                                   status ? status + o : nullptr, s);
                      });
}

}  // extern "C"

// ---- imageToRemoteImagingLocus / computeGroundPartials ---------------------------------------
namespace ub200 {

cudaError_t launchRemoteLocus(Model &m, long n, const double *line, const double *samp,
                                     double prec, const LocusOut &o, cudaStream_t stream) {
  switch (m.kind) {
    case USGSCSM_B200_MODEL_LINE_SCANNER:
      return launchLsRemoteLocus(m.ls, m.smCount, n, line, samp, o, stream);
    case USGSCSM_B200_MODEL_PUSH_FRAME:
      return launchPfRemoteLocus(m.ls, m.smCount, n, line, samp, o, stream);
    case USGSCSM_B200_MODEL_FRAME:
      return launchFrRemoteLocus(m.fr, m.smCount, n, line, samp, o, stream);
    case USGSCSM_B200_MODEL_SAR:
      return launchSarRemoteLocus(m.sar, m.smCount, n, line, samp, prec, o, stream);
    default:
      return cudaErrorNotSupported;
  }
}

cudaError_t launchProximateLocus(Model &m, long n, const double *line, const double *samp,
                                        const double *x, const double *y, const double *z,
                                        double prec, const LocusOut &o, cudaStream_t stream) {
  switch (m.kind) {
    case USGSCSM_B200_MODEL_LINE_SCANNER:
      return launchLsProximateLocus(m.ls, m.smCount, n, line, samp, x, y, z, prec, o, stream);
    case USGSCSM_B200_MODEL_PUSH_FRAME:
      return launchPfProximateLocus(m.ls, m.smCount, n, line, samp, x, y, z, prec, o, stream);
    case USGSCSM_B200_MODEL_FRAME:
      return launchFrProximateLocus(m.fr, m.smCount, n, line, samp, x, y, z, o, stream);
    case USGSCSM_B200_MODEL_SAR:
      return launchSarProximateLocus(m.sar, m.smCount, n, line, samp, x, y, z, o, stream);
    default:
      return cudaErrorNotSupported;
  }
}

cudaError_t launchGroundPartials(Model &m, long n, const double *x, const double *y,
                                        const double *z, double *partials, uint32_t *status,
                                        cudaStream_t stream) {
  switch (m.kind) {
    case USGSCSM_B200_MODEL_LINE_SCANNER:
      return launchGroundPartialsLs(m.ls, m.smCount, n, x, y, z, partials, status, stream);
    case USGSCSM_B200_MODEL_PUSH_FRAME:
      return launchGroundPartialsPf(m.ls, m.smCount, n, x, y, z, partials, status, stream);
    case USGSCSM_B200_MODEL_FRAME:
      return launchGroundPartialsFr(m.fr, m.smCount, n, x, y, z, partials, status, stream);
    case USGSCSM_B200_MODEL_SAR:
      return launchGroundPartialsSar(m.sar, m.smCount, n, x, y, z, partials, status, stream);
    default:
      return cudaErrorNotSupported;
  }
}

}  // namespace ub200

extern "C" {

int usgscsm_b200_image_to_remote_imaging_locus_dev(usgscsm_b200_model *m, int64_t n,
                                                   const double *line, const double *samp,
                                                   double prec, double *px, double *py, double *pz,
                                                   double *dx, double *dy, double *dz,
                                                   double *achieved, uint32_t *status, void *stream) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!line || !samp || !px || !py || !pz || !dx || !dy || !dz)
    return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    LocusOut o{px, py, pz, dx, dy, dz, achieved, status};
    UB_CUDA(launchRemoteLocus(*m, (long)n, line, samp, prec, o, static_cast<cudaStream_t>(stream)));
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_ground_partials_dev(usgscsm_b200_model *m, int64_t n, const double *x,
                                     const double *y, const double *z, double *partials,
                                     uint32_t *status, void *stream) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!x || !y || !z || !partials) return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    UB_CUDA(launchGroundPartials(*m, (long)n, x, y, z, partials, status,
                                 static_cast<cudaStream_t>(stream)));
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_image_to_proximate_imaging_locus_dev(
    usgscsm_b200_model *m, int64_t n, const double *line, const double *samp, const double *x,
    const double *y, const double *z, double prec, double *px, double *py, double *pz, double *dx,
    double *dy, double *dz, double *achieved, uint32_t *status, void *stream) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!line || !samp || !x || !y || !z || !px || !py || !pz || !dx || !dy || !dz)
    return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    LocusOut o{px, py, pz, dx, dy, dz, achieved, status};
    UB_CUDA(launchProximateLocus(*m, (long)n, line, samp, x, y, z, prec, o,
                                 static_cast<cudaStream_t>(stream)));
    return USGSCSM_B200_OK;
  });
}

}  // extern "C"

namespace ub200 {

// computeSensorPartials(index, ground_pt, prec, ach, warn) LS:1336-1396 over n points: the
// search with the model as it is (skipped when the caller supplies the image points, the
// 6-argument overload), the search with parameter `index` moved by m_gsd, forward difference.
// DELTA of the scalar call: m_gsd (LS:1386, PF:1153), 1.0 or 0.01 for the quaternion
// parameters (FR:853-859), m_scaledPixelWidth (SAR:1360)
static double sensorPartialDelta(const usgscsm_b200_model *m, int index) {
  switch (m->kind) {
    case USGSCSM_B200_MODEL_FRAME: return (index > 2 && index < 7) ? 0.01 : 1.0;
    case USGSCSM_B200_MODEL_SAR: return m->sar.scaledPixelWidth;
    default: return m->ls.gsd;
  }
}

// groundToImage with parameter `index` moved by delta.  *pfTable: scratch the caller frees on
// the stream afterwards (per-framelet orientations of the adjusted push-frame model).
static cudaError_t launchAdjustedGroundToImage(usgscsm_b200_model *m, int index, double delta,
                                               long n, const double *x, const double *y,
                                               const double *z, double prec, double *l1,
                                               double *s1, double *achieved, uint32_t *st1,
                                               cudaStream_t stream, double **pfTable) {
  switch (m->kind) {
    case USGSCSM_B200_MODEL_LINE_SCANNER:
      return launchLsGroundToImage(adjustedLineScanner(*m, index, delta), m->smCount, n, x, y, z,
                                   prec, l1, s1, achieved, st1, stream);
    case USGSCSM_B200_MODEL_PUSH_FRAME: {
      LsState adj = adjustedPushFrame(*m, index, delta);
      if (m->ls.frameletOrient) {  // per-framelet orientations of the adjusted model
        cudaError_t e = cudaMallocAsync(
            pfTable, (size_t)(adj.numFramelets + 2) * kFrameletOrientDoubles * sizeof(double), stream);
        if (e == cudaSuccess) e = launchPfFrameletOrientations(adj, *pfTable, stream);
        if (e != cudaSuccess) return e;
        adj.frameletOrient = *pfTable;
      }
      return launchPfGroundToImage(adj, m->smCount, n, x, y, z, prec, l1, s1, achieved, st1, stream);
    }
    case USGSCSM_B200_MODEL_FRAME:
      return launchFrGroundToImage(adjustedFrame(*m, index, delta), m->smCount, n, x, y, z, prec,
                                   l1, s1, achieved, st1, stream);
    default:
      return launchSarGroundToImage(adjustedSar(*m, index, delta), m->smCount, n, x, y, z, prec, l1,
                                    s1, achieved, st1, stream);
  }
}

int sensorPartialsDev(usgscsm_b200_model *m, int index, long n, const double *x,
                             const double *y, const double *z, const double *img_line,
                             const double *img_samp, double prec, double *dline, double *dsamp,
                             double *achieved, uint32_t *status, cudaStream_t stream) {
  const double delta = sensorPartialDelta(m, index);
  const bool haveImage = img_line && img_samp;
  // scratch: l0, s0 (unless supplied), l1, s1, st0, st1 (unless the caller wants status)
  double *scratch = nullptr, *pfTable = nullptr;
  uint32_t *stScratch = nullptr;
  UB_CUDA(cudaMallocAsync(&scratch, (size_t)n * 4 * sizeof(double), stream));
  UB_CUDA(cudaMallocAsync(&stScratch, (size_t)n * 2 * sizeof(uint32_t), stream));
  double *l0 = scratch, *s0 = scratch + n, *l1 = scratch + 2 * n, *s1 = scratch + 3 * n;
  uint32_t *st0 = stScratch, *st1 = status ? status : stScratch + n;
  cudaError_t e = cudaSuccess;
  if (!haveImage)
    e = launchGroundToImage(*m, n, x, y, z, prec, l0, s0, nullptr, st0, stream);
  if (e == cudaSuccess)
    e = launchAdjustedGroundToImage(m, index, delta, n, x, y, z, prec, l1, s1, achieved, st1,
                                    stream, &pfTable);
  if (e == cudaSuccess)
    e = launchSensorPartialsCombine(n, haveImage ? img_line : l0, haveImage ? img_samp : s0,
                                    haveImage ? nullptr : st0, l1, s1, delta, dline, dsamp,
                                    achieved, st1, stream);
  cudaFreeAsync(scratch, stream);
  cudaFreeAsync(stScratch, stream);
  if (pfTable) cudaFreeAsync(pfTable, stream);
  UB_CUDA(e);
  return USGSCSM_B200_OK;
}

// computeAllSensorPartials (csm/RasterGM.cpp:52-140 through LS:1413-1455, PF:1170-1201,
// FR:894-940): the image point once (unless supplied), then one adjusted search per requested
// parameter, all against that image point; achieved precision = the largest over the
// parameters starting from 0, warnings of every search accumulate.  partials: per point,
// per requested parameter, (dline, dsamp).
int allSensorPartialsDev(usgscsm_b200_model *m, int n_indices, const int *indices, long n,
                         const double *x, const double *y, const double *z,
                         const double *img_line, const double *img_samp, double prec,
                         double *partials, double *achieved, uint32_t *status,
                         cudaStream_t stream) {
  const bool haveImage = img_line && img_samp;
  if (m->kind == USGSCSM_B200_MODEL_LINE_SCANNER) {
    // one kernel: the searches share the seed and the interpolations of their first two evaluations
    cudaError_t fe = cudaSuccess;
    if (launchLsAllSensorPartials(m->ls, m->smCount, n_indices, indices, n, x, y, z,
                                  haveImage ? img_line : nullptr, haveImage ? img_samp : nullptr, prec,
                                  partials, achieved, status, stream, fe)) {
      UB_CUDA(fe);
      return USGSCSM_B200_OK;
    }
  }
  double *scratch = nullptr;
  uint32_t *stScratch = nullptr;
  std::vector<double *> pfTables;
  UB_CUDA(cudaMallocAsync(&scratch, (size_t)n * 5 * sizeof(double), stream));
  UB_CUDA(cudaMallocAsync(&stScratch, (size_t)n * 3 * sizeof(uint32_t), stream));
  double *l0 = scratch, *s0 = scratch + n, *l1 = scratch + 2 * n, *s1 = scratch + 3 * n,
         *achP = scratch + 4 * n;
  uint32_t *st0 = stScratch, *stP = stScratch + n, *acc = status ? status : stScratch + 2 * n;
  cudaError_t e = cudaSuccess;
  if (!haveImage)
    e = launchGroundToImage(*m, n, x, y, z, prec, l0, s0, nullptr, st0, stream);
  for (int k = 0; k < n_indices && e == cudaSuccess; k++) {
    const int index = indices ? indices[k] : k;
    const double delta = sensorPartialDelta(m, index);
    if (m->kind == USGSCSM_B200_MODEL_LINE_SCANNER) {
      // the search writes its forward difference itself (no separate step kernel)
      e = launchLsSensorPartialStep(adjustedLineScanner(*m, index, delta), m->smCount, n, x, y, z, prec,
                                    haveImage ? img_line : l0, haveImage ? img_samp : s0,
                                    haveImage ? nullptr : st0, delta, partials, n_indices, k, achieved,
                                    acc, stream);
      continue;
    }
    double *pfTable = nullptr;
    e = launchAdjustedGroundToImage(m, index, delta, n, x, y, z, prec, l1, s1, achP, stP, stream,
                                    &pfTable);
    if (pfTable) pfTables.push_back(pfTable);
    if (e == cudaSuccess)
      e = launchSensorPartialsStep(n, haveImage ? img_line : l0, haveImage ? img_samp : s0,
                                   haveImage ? nullptr : st0, l1, s1, achP, stP, delta, partials,
                                   n_indices, k, achieved, acc, stream);
  }
  cudaFreeAsync(scratch, stream);
  cudaFreeAsync(stScratch, stream);
  for (double *t : pfTables) cudaFreeAsync(t, stream);
  UB_CUDA(e);
  return USGSCSM_B200_OK;
}

int allSensorPartialsCheck(usgscsm_b200_model *m, int n_indices, const int *indices, int64_t n,
                           const void *x, const void *y, const void *z, const void *partials) {
  if (!m || n < 0 || n_indices < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (!indices && n_indices != (int)m->params.size())
    return fail(USGSCSM_B200_EINVAL, "without an index list n_indices must be the parameter count");
  for (int k = 0; indices && k < n_indices; k++)
    if (indices[k] < 0 || indices[k] >= (int)m->params.size())
      return fail(USGSCSM_B200_EINVAL, "parameter index out of range");
  if (n > 0 && n_indices > 0 && (!x || !y || !z || !partials))
    return fail(USGSCSM_B200_EINVAL, "null array");
  return USGSCSM_B200_OK;
}

int sensorPartialsCheck(usgscsm_b200_model *m, int index, int64_t n, const void *x,
                               const void *y, const void *z, const void *dline,
                               const void *dsamp) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (index < 0 || index >= (int)m->params.size())
    return fail(USGSCSM_B200_EINVAL, "parameter index out of range");
  if (n > 0 && (!x || !y || !z || !dline || !dsamp)) return fail(USGSCSM_B200_EINVAL, "null array");
  return USGSCSM_B200_OK;
}

}  // namespace ub200

extern "C" {

int usgscsm_b200_sensor_partials_dev(usgscsm_b200_model *m, int index, int64_t n, const double *x,
                                     const double *y, const double *z, const double *img_line,
                                     const double *img_samp, double prec, double *dline,
                                     double *dsamp, double *achieved, uint32_t *status,
                                     void *stream) {
  int c = sensorPartialsCheck(m, index, n, x, y, z, dline, dsamp);
  if (c != USGSCSM_B200_OK || n == 0) return c;
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    return sensorPartialsDev(m, index, (long)n, x, y, z, img_line, img_samp, prec, dline, dsamp,
                             achieved, status, static_cast<cudaStream_t>(stream));
  });
}

// getImageTime / getSensorPosition / getSensorVelocity of an image point, for arrays
// (LS:1543-1653, PF:1258-1320, FR:665-765, SAR:1113-1238).
int usgscsm_b200_image_sensor_state_dev(usgscsm_b200_model *m, int64_t n, const double *line,
                                        const double *samp, double *time, double *px, double *py,
                                        double *pz, double *vx, double *vy, double *vz,
                                        uint32_t *status, void *stream) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!line || !samp) return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    SensorStateOut o{time, px, py, pz, vx, vy, vz, status, 0};
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (m->kind) {
      case USGSCSM_B200_MODEL_LINE_SCANNER: UB_CUDA(launchLsSensorState(m->ls, (long)n, line, o, s)); break;
      case USGSCSM_B200_MODEL_PUSH_FRAME: UB_CUDA(launchPfSensorState(m->ls, (long)n, line, o, s)); break;
      case USGSCSM_B200_MODEL_FRAME: UB_CUDA(launchFrSensorState(m->fr, (long)n, line, samp, o, s)); break;
      default: UB_CUDA(launchSarSensorState(m->sar, (long)n, line, o, s)); break;
    }
    return USGSCSM_B200_OK;
  });
}

// getSensorPosition(double time) / getSensorVelocity(double time) for arrays of image times
// (LS:1599-1653, PF:1281-1320, FR:707-795, SAR:1160-1260).
int usgscsm_b200_sensor_state_at_time_dev(usgscsm_b200_model *m, int64_t n, const double *time,
                                          double *px, double *py, double *pz, double *vx,
                                          double *vy, double *vz, uint32_t *status, void *stream) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!time) return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    SensorStateOut o{nullptr, px, py, pz, vx, vy, vz, status, 1};
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    switch (m->kind) {
      case USGSCSM_B200_MODEL_LINE_SCANNER: UB_CUDA(launchLsSensorState(m->ls, (long)n, time, o, s)); break;
      case USGSCSM_B200_MODEL_PUSH_FRAME: UB_CUDA(launchPfSensorState(m->ls, (long)n, time, o, s)); break;
      case USGSCSM_B200_MODEL_FRAME: UB_CUDA(launchFrSensorState(m->fr, (long)n, time, time, o, s)); break;
      default: UB_CUDA(launchSarSensorState(m->sar, (long)n, time, o, s)); break;
    }
    return USGSCSM_B200_OK;
  });
}

int usgscsm_b200_all_sensor_partials_dev(usgscsm_b200_model *m, int n_indices, const int *indices,
                                         int64_t n, const double *x, const double *y,
                                         const double *z, const double *img_line,
                                         const double *img_samp, double prec, double *partials,
                                         double *achieved, uint32_t *status, void *stream) {
  int c = allSensorPartialsCheck(m, n_indices, indices, n, x, y, z, partials);
  if (c != USGSCSM_B200_OK || n == 0 || n_indices == 0) return c;
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    return allSensorPartialsDev(m, n_indices, indices, (long)n, x, y, z, img_line, img_samp, prec,
                                partials, achieved, status, static_cast<cudaStream_t>(stream));
  });
}

// csm::SettableEllipsoid::setEllipsoid / getEllipsoid (LS:2116-2119, PF:1663-1666,
// FR:2235-2240, SAR:1884-1887): replaces the two radii, nothing else is re-derived.
int usgscsm_b200_model_set_ellipsoid(usgscsm_b200_model *m, double semi_major, double semi_minor) {
  if (!m) return fail(USGSCSM_B200_EINVAL, "null argument");
  Ellipsoid &e = m->kind == USGSCSM_B200_MODEL_FRAME ? m->fr.ell
                 : m->kind == USGSCSM_B200_MODEL_SAR ? m->sar.ell : m->ls.ell;
  setEllipsoidRadii(e, semi_major, semi_minor);
  return USGSCSM_B200_OK;
}

int usgscsm_b200_model_get_ellipsoid(const usgscsm_b200_model *m, double *semi_major,
                                     double *semi_minor) {
  if (!m || !semi_major || !semi_minor) return fail(USGSCSM_B200_EINVAL, "null argument");
  const Ellipsoid &e = m->kind == USGSCSM_B200_MODEL_FRAME ? m->fr.ell
                       : m->kind == USGSCSM_B200_MODEL_SAR ? m->sar.ell : m->ls.ell;
  *semi_major = e.majorAxis;
  *semi_minor = e.minorAxis;
  return USGSCSM_B200_OK;
}

}  // extern "C"

// ---- building blocks ------------------------------------------------------------------------
namespace {
struct DevArrays {  // a few device arrays freed on scope exit
  std::vector<void *> ptrs;
  ~DevArrays() {
    for (void *p : ptrs) cudaFree(p);
  }
  template <class T>
  T *alloc(size_t count) {
    void *p = nullptr;
    UB_CUDA(cudaMalloc(&p, (count ? count : 1) * sizeof(T)));
    ptrs.push_back(p);
    return static_cast<T *>(p);
  }
  template <class T>
  T *upload(const T *host, size_t count) {
    T *d = alloc<T>(count);
    if (count) UB_CUDA(cudaMemcpy(d, host, count * sizeof(T), cudaMemcpyHostToDevice));
    return d;
  }
};

int distortionCall(bool apply, int device, int type, const double *coeffs, int ncoeffs,
                   double focal, double precision, int64_t n, const double *inX,
                   const double *inY, double *outX, double *outY, uint32_t *status) {
  if (n < 0 || ncoeffs < 0 || (ncoeffs && !coeffs)) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (!inX || !inY || !outX || !outY) return fail(USGSCSM_B200_EINVAL, "null array");
  if (device < 0 || device >= deviceCount())
    return fail(USGSCSM_B200_ENODEVICE, "no usable CUDA device with that index");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(device));
    checkDistortionCoeffs(type, (size_t)ncoeffs);
    FocalPlane fp{};
    fp.distortionType = type;
    fp.numDistCoeffs = ncoeffs;
    for (int i = 0; i < ncoeffs; i++) fp.distCoeffs[i] = coeffs[i];
    fp.focalLength = focal;
    fp.detectorLineSumming = fp.detectorSampleSumming = 1.0;
    DevArrays d;
    const double *dx = d.upload(inX, (size_t)n), *dy = d.upload(inY, (size_t)n);
    double *ox = d.alloc<double>((size_t)n), *oy = d.alloc<double>((size_t)n);
    uint32_t *st = status ? d.alloc<uint32_t>((size_t)n) : nullptr;
    UB_CUDA(launchDistortion(apply, fp, precision, (long)n, dx, dy, ox, oy, st, nullptr));
    UB_CUDA(cudaMemcpy(outX, ox, n * sizeof(double), cudaMemcpyDeviceToHost));
    UB_CUDA(cudaMemcpy(outY, oy, n * sizeof(double), cudaMemcpyDeviceToHost));
    if (status) UB_CUDA(cudaMemcpy(status, st, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    return USGSCSM_B200_OK;
  });
}
}  // namespace

extern "C" {

int usgscsm_b200_apply_distortion(int device, int type, const double *coeffs, int ncoeffs,
                                  double focal_length, double desired_precision, int64_t n,
                                  const double *ux, const double *uy, double *dx, double *dy,
                                  uint32_t *status) {
  return distortionCall(true, device, type, coeffs, ncoeffs, focal_length, desired_precision, n,
                        ux, uy, dx, dy, status);
}

int usgscsm_b200_remove_distortion(int device, int type, const double *coeffs, int ncoeffs,
                                   double focal_length, double tolerance, int64_t n,
                                   const double *dx, const double *dy, double *ux, double *uy,
                                   uint32_t *status) {
  return distortionCall(false, device, type, coeffs, ncoeffs, focal_length, tolerance, n, dx, dy,
                        ux, uy, status);
}

int usgscsm_b200_lagrange_interp(int device, int num_time, const double *values, double start_time,
                                 double del_time, int vector_length, int order, int64_t n,
                                 const double *times, double *out) {
  if (n < 0 || !values || vector_length < 1 || vector_length > 4)
    return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (num_time < 2)  // U:414-419 throws csm::Error(INDEX_OUT_OF_RANGE)
    return fail(USGSCSM_B200_EINVAL, "Not enough data to interpolate");
  if (n == 0) return USGSCSM_B200_OK;
  if (!times || !out) return fail(USGSCSM_B200_EINVAL, "null array");
  if (device < 0 || device >= deviceCount())
    return fail(USGSCSM_B200_ENODEVICE, "no usable CUDA device with that index");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(device));
    DevArrays d;
    UniformTable tab{d.upload(values, (size_t)num_time * vector_length), num_time, order,
                     start_time, del_time, 0.0};
    fillTableSteps({&tab});
    LagrangeDivisors lag;
    std::vector<Divisor *> ds;
    appendLagrangeDivisors(lag, ds);
    fillDivisors(ds);
    const double *dt = d.upload(times, (size_t)n);
    double *dout = d.alloc<double>((size_t)n * vector_length);
    UB_CUDA(launchLagrange(tab, lag, vector_length, (long)n, dt, dout, nullptr));
    UB_CUDA(cudaMemcpy(out, dout, n * vector_length * sizeof(double), cudaMemcpyDeviceToHost));
    return USGSCSM_B200_OK;
  });
}

// ---- the reference's remaining free functions, for known-answer replay ------------------------
}  // extern "C"
namespace ub200 {
void fillFocalPlaneInverse(FocalPlane &fp);  // model_ls.cu
}
namespace {
FocalPlane affineFocalPlane(const double a[12]) {
  FocalPlane fp{};
  for (int i = 0; i < 3; i++) {
    fp.iTransS[i] = a[i];
    fp.iTransL[i] = a[3 + i];
  }
  fp.detectorSampleOrigin = a[6];
  fp.detectorLineOrigin = a[7];
  fp.detectorSampleSumming = a[8];
  fp.detectorLineSumming = a[9];
  fp.startingDetectorSample = a[10];
  fp.startingDetectorLine = a[11];
  fillFocalPlaneInverse(fp);
  fp.lineSumming.b = fp.detectorLineSumming;
  fp.sampleSumming.b = fp.detectorSampleSumming;
  fillDivisors({&fp.lineSumming, &fp.sampleSumming});
  return fp;
}

// n-element host arrays in, n * outWidth-element host arrays out, through one small kernel
template <class Launch>
int katCall(int device, int64_t n, Launch &&launch) {
  if (n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (n == 0) return USGSCSM_B200_OK;
  if (device < 0 || device >= deviceCount())
    return fail(USGSCSM_B200_ENODEVICE, "no usable CUDA device with that index");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(device));
    DevArrays d;
    launch(d);
    return USGSCSM_B200_OK;
  });
}
template <class T>
void download(T *host, const T *dev, size_t count) {
  if (host) UB_CUDA(cudaMemcpy(host, dev, count * sizeof(T), cudaMemcpyDeviceToHost));
}
}  // namespace
extern "C" {

int usgscsm_b200_focal_plane_coordinates(int device, const double affine12[12], int64_t n,
                                         const double *line, const double *sample, double *focal_x,
                                         double *focal_y) {
  if (!affine12 || (n > 0 && (!line || !sample || !focal_x || !focal_y)))
    return fail(USGSCSM_B200_EINVAL, "null argument");
  return katCall(device, n, [&](DevArrays &d) {
    const FocalPlane fp = affineFocalPlane(affine12);
    const double *a = d.upload(line, (size_t)n), *b = d.upload(sample, (size_t)n);
    double *oa = d.alloc<double>((size_t)n), *ob = d.alloc<double>((size_t)n);
    UB_CUDA(launchFocalPlane(fp, 0, (long)n, a, b, oa, ob, nullptr));
    download(focal_x, oa, (size_t)n);
    download(focal_y, ob, (size_t)n);
  });
}

int usgscsm_b200_compute_pixel(int device, const double affine12[12], int64_t n, const double *focal_x,
                               const double *focal_y, double *line, double *sample) {
  if (!affine12 || (n > 0 && (!line || !sample || !focal_x || !focal_y)))
    return fail(USGSCSM_B200_EINVAL, "null argument");
  return katCall(device, n, [&](DevArrays &d) {
    const FocalPlane fp = affineFocalPlane(affine12);
    const double *a = d.upload(focal_x, (size_t)n), *b = d.upload(focal_y, (size_t)n);
    double *oa = d.alloc<double>((size_t)n), *ob = d.alloc<double>((size_t)n);
    UB_CUDA(launchFocalPlane(fp, 1, (long)n, a, b, oa, ob, nullptr));
    download(line, oa, (size_t)n);
    download(sample, ob, (size_t)n);
  });
}

int usgscsm_b200_camera_look_vector(int device, double z_direction, double focal_length, int64_t n,
                                    const double *focal_x, const double *focal_y, double *look) {
  if (n > 0 && (!focal_x || !focal_y || !look)) return fail(USGSCSM_B200_EINVAL, "null argument");
  return katCall(device, n, [&](DevArrays &d) {
    const double *a = d.upload(focal_x, (size_t)n), *b = d.upload(focal_y, (size_t)n);
    double *o = d.alloc<double>((size_t)n * 3);
    UB_CUDA(launchLookVector(z_direction, focal_length, (long)n, a, b, o, nullptr));
    download(look, o, (size_t)n * 3);
  });
}

int usgscsm_b200_jitter(int device, int remove, int num_coeffs, const double *line_jitter,
                        const double *sample_jitter, int num_line_times, const double *line_times,
                        double tolerance, int max_iterations, int64_t n, const double *line,
                        const double *sample, double *out_line, double *out_sample) {
  if (num_coeffs < 0 || num_line_times < 1 || !line_times || (num_coeffs && (!line_jitter || !sample_jitter)) ||
      (n > 0 && (!line || !sample || !out_line || !out_sample)))
    return fail(USGSCSM_B200_EINVAL, "bad argument");
  return katCall(device, n, [&](DevArrays &d) {
    FrameState S{};
    S.numJitterCoeffs = num_coeffs;
    S.numLineTimes = num_line_times;
    S.lineJitter = d.upload(line_jitter, (size_t)num_coeffs);
    S.sampleJitter = d.upload(sample_jitter, (size_t)num_coeffs);
    S.lineTimes = d.upload(line_times, (size_t)num_line_times);
    const double *a = d.upload(line, (size_t)n), *b = d.upload(sample, (size_t)n);
    double *oa = d.alloc<double>((size_t)n), *ob = d.alloc<double>((size_t)n);
    UB_CUDA(launchJitter(S, remove, tolerance, max_iterations, (long)n, a, b, oa, ob, nullptr));
    download(out_line, oa, (size_t)n);
    download(out_sample, ob, (size_t)n);
  });
}

int usgscsm_b200_polynomial(int device, const double coeffs4[4], int64_t n, const double *x,
                            double *value, double *derivative) {
  if (!coeffs4 || (n > 0 && !x)) return fail(USGSCSM_B200_EINVAL, "null argument");
  return katCall(device, n, [&](DevArrays &d) {
    const double *a = d.upload(x, (size_t)n);
    double *o0 = value ? d.alloc<double>((size_t)n) : nullptr;
    double *o1 = derivative ? d.alloc<double>((size_t)n) : nullptr;
    UB_CUDA(launchCubic(coeffs4, 0, 0.0, (long)n, a, nullptr, o0, o1, nullptr, nullptr));
    download(value, o0, (size_t)n);
    download(derivative, o1, (size_t)n);
  });
}

int usgscsm_b200_polynomial_root(int device, const double coeffs4[4], double threshold, int64_t n,
                                 const double *guess, double *root, uint32_t *status) {
  if (!coeffs4 || (n > 0 && (!guess || !root))) return fail(USGSCSM_B200_EINVAL, "null argument");
  return katCall(device, n, [&](DevArrays &d) {
    const double *a = d.upload(guess, (size_t)n);
    double *o0 = d.alloc<double>((size_t)n);
    uint32_t *st = status ? d.alloc<uint32_t>((size_t)n) : nullptr;
    UB_CUDA(launchCubic(coeffs4, 1, threshold, (long)n, a, nullptr, o0, nullptr, st, nullptr));
    download(root, o0, (size_t)n);
    download(status, st, (size_t)n);
  });
}

int usgscsm_b200_brent_root(int device, const double coeffs4[4], double epsilon, int64_t n,
                            const double *lower, const double *upper, double *root,
                            uint32_t *status) {
  if (!coeffs4 || (n > 0 && (!lower || !upper || !root))) return fail(USGSCSM_B200_EINVAL, "null argument");
  return katCall(device, n, [&](DevArrays &d) {
    const double *a = d.upload(lower, (size_t)n), *b = d.upload(upper, (size_t)n);
    double *o0 = d.alloc<double>((size_t)n);
    uint32_t *st = status ? d.alloc<uint32_t>((size_t)n) : nullptr;
    UB_CUDA(launchCubic(coeffs4, 2, epsilon, (long)n, a, b, o0, nullptr, st, nullptr));
    download(root, o0, (size_t)n);
    download(status, st, (size_t)n);
  });
}

int usgscsm_b200_selftest_division(int device, int64_t pairs, uint64_t seed, uint64_t *mismatches) {
  if (!mismatches || pairs < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (device < 0 || device >= deviceCount())
    return fail(USGSCSM_B200_ENODEVICE, "no usable CUDA device with that index");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(device));
    DevArrays d;
    unsigned long long *dm = d.alloc<unsigned long long>(1);
    UB_CUDA(cudaMemset(dm, 0, sizeof(unsigned long long)));
    const int blocks = 148 * 8;
    long perThread = (long)((pairs + (int64_t)blocks * 256 - 1) / ((int64_t)blocks * 256));
    UB_CUDA(launchDivisionSelfTest((unsigned long long)seed, blocks, perThread, dm, nullptr));
    unsigned long long h = 0;
    UB_CUDA(cudaMemcpy(&h, dm, sizeof(h), cudaMemcpyDeviceToHost));
    *mismatches = h;
    return USGSCSM_B200_OK;
  });
}

}  // extern "C"
