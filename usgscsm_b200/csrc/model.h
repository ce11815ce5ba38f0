// model.h -- host-side model object behind the C-ABI handle (usgscsm_b200_model).
//
// Mirrors what a reference model object owns on the hot path: the members read by
// groundToImage / imageToGround (host copies, for re-packing after setParameterValue)
// and, new here, the packed device state + device tables + streams/staging buffers.
#ifndef USGSCSM_B200_MODEL_H_
#define USGSCSM_B200_MODEL_H_

#include <cuda_runtime.h>

#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "json_min.h"
#include "kernels.h"
#include "sensor_state.h"

namespace ub200 {

struct CudaError : std::runtime_error {
  explicit CudaError(const std::string &what) : std::runtime_error(what) {}
};
struct StateError : std::runtime_error {
  explicit StateError(const std::string &what) : std::runtime_error(what) {}
};

#define UB_CUDA(expr)                                                                     \
  do {                                                                                    \
    cudaError_t e_ = (expr);                                                              \
    if (e_ != cudaSuccess)                                                                \
      throw ::ub200::CudaError(std::string(#expr) + ": " + cudaGetErrorString(e_));       \
  } while (0)

// Device allocation owned by a model.
struct DeviceBuffer {
  void *ptr = nullptr;
  size_t bytes = 0;
};

// One slot of the chunked host-pointer pipeline (capi_host.cu): device scratch for one chunk's
// inputs, outputs and status words, a stream, and the event that marks the chunk done.
struct PipelineSlot {
  cudaStream_t stream = nullptr;
  cudaEvent_t done = nullptr;
  double *scratch = nullptr;
  long scratchDoubles = 0;
  // page-locked host buffer of the same layout, for callers whose arrays are pageable
  void *stage = nullptr;
  long stageBytes = 0;
};

constexpr int kMaxPipelineSlots = 16;
constexpr int kDefaultPipelineSlots = 3;
constexpr long kDefaultPipelineChunk = 1L << 22;  // 4 Mi points per chunk (~200 MB of I/O)
// USGSCSM_B200_PIPELINE_SLOTS / USGSCSM_B200_PIPELINE_CHUNK override the defaults (capi.cu)
int pipelineSlots();
long pipelineChunk();

struct Model {
  int kind = 0;
  int device = 0;
  int smCount = 0;
  std::string modelName;

  // ---- host copies of the members the per-point math reads ---------------------------
  // line-scan / push-frame
  LsState ls{};  // scalars valid on host; pointer members hold DEVICE addresses
  std::vector<double> intTimeLines, intTimeStartTimes, intTimes;
  std::vector<double> positions, velocities, quaternions;
  double referencePoint[3] = {0, 0, 0};
  int platformFlag = 1;
  int numQuaternionDoubles = 0;  // m_numQuaternions (counts doubles, LS:2145)
  // frame
  FrameState fr{};
  std::vector<double> lineJitter, sampleJitter, lineTimes;
  // SAR
  SarState sar{};
  std::vector<double> scaleCoeffs, scaleTimes;

  std::vector<double> params;  // m_currentParameterValue (raw, without the +0.0)
  json::Value stateTree;       // the state as loaded; getModelState re-serialises it
  std::vector<double> covariance;  // m_covariance, numParams x numParams (error propagation only)
  // getIlluminationDirection only (illumination.cu)
  std::vector<double> sunPosition, sunVelocity;
  UniformTable sunTable{};
  LagrangeDivisors sunLag{};
  bool hasSunTable = false;

  // ---- device side ---------------------------------------------------------------------
  std::vector<DeviceBuffer> owned;
  double *frameletOrient = nullptr;  // push-frame: device table behind ls.frameletOrient
  PipelineSlot slots[kMaxPipelineSlots];
  // Host-pointer entry points share the slots' scratch and streams: one call at a time per model
  // (capi_host.cu runPipelined).  Concurrent host threads on ONE handle are therefore safe and
  // serialised; threads on different handles run concurrently.
  std::mutex hostCallMutex;
  // Row / column trigonometry tables of the DEM windows last used with this model (capi_geo.cu):
  // a repeated call on the same window (an orthorectifier re-projecting tile after tile, the
  // bench loop) re-uses the device copy instead of rebuilding it on the host.
  struct GeoTables {
    std::string key;
    double *dev = nullptr;
    size_t rowDoubles = 0;
  };
  std::vector<GeoTables> geoTables;
  std::mutex geoTablesMutex;

  ~Model();
  const double *upload(const std::vector<double> &v);
  void ensureScratch(int s, long doubles);
  void ensureStage(int s, long bytes);
};

// model_ls.cu
void loadLineScanner(Model &m, const class StateReader &st);
void repackLineScanner(Model &m);  // after a parameter / projective-seed change
void fillDefaultCovariance(std::vector<double> &cov, double gsd);  // updateState LS:703-734
LsState adjustedLineScanner(const Model &m, int index, double delta);
// model_pf.cu
void loadPushFrame(Model &m, const class StateReader &st);
void repackPushFrame(Model &m);
LsState adjustedPushFrame(const Model &m, int index, double delta);  // after a parameter change: params + framelet orientations
// divisors.cu: reciprocals of constant divisors, evaluated on the device
void fillDivisors(const std::vector<Divisor *> &divisors);
void appendLagrangeDivisors(LagrangeDivisors &L, std::vector<Divisor *> &out);
void fillTableSteps(const std::vector<UniformTable *> &tables);
// model_fr.cu
void loadFrame(Model &m, const class StateReader &st);
void repackFrame(Model &m);
FrameState adjustedFrame(const Model &m, int index, double delta);
// model_sar.cu
void loadSar(Model &m, const class StateReader &st);
void repackSar(Model &m);
SarState adjustedSar(const Model &m, int index, double delta);
// shared readers (model_ls.cu)
void readFocalPlane(FocalPlane &fp, const class StateReader &st, bool lineScanTypes);
void readEllipsoid(Ellipsoid &e, const class StateReader &st);
void setEllipsoidRadii(Ellipsoid &e, double a, double b);
void readOrientation(Model &m, const class StateReader &st);

// illumination.cu
void loadSun(Model &m, const class StateReader &st);

// projective_fit.cpp
// Least-squares fit of the 14-coefficient 3D->2D projective map (the reference's
// usgscsm::computeBestFitProjectiveTransform, src/EigenUtilities.cpp:14-66).
void fitProjectiveTransform(int numPts, const double *line, const double *samp,
                            const double *x, const double *y, const double *z,
                            double coeffs[14]);

}  // namespace ub200

// The opaque C handle is the Model itself.
struct usgscsm_b200_model : ub200::Model {};

#endif
