// covar.cu -- error propagation around the path: groundToImage(csm::EcefCoordCovar) and
// imageToGround(csm::ImageCoordCovar, height, heightVariance) of the line-scan, push-frame and
// SAR models (LS:910-964 / 1075-1157, PF:784-840 / 909-990, SAR:709-758 / 864-940), batched.
//
// The scalar calls are compositions of the per-point searches: the plain projection, ground
// partials (four searches), determineSensorCovarianceInImageSpace (LS:1025-1054: all sensor
// partials at the default precision 0.001, then sum_ij p_i C_ij p_j), and three more
// imageToGround for the image -> ground direction.  Here the searches are the existing batched
// kernels and one small kernel per direction does the matrix algebra in the reference's
// operation order.  The frame model's overloads are stubs in the reference (FR:234-278,
// 374-395: "incomplete implementation"); they are refused here.
#include <vector>

#include "capi_common.h"
#include "kernels.h"

namespace ub200 {

namespace {

constexpr int kMaxParams = 16;
constexpr long kCovarBatch = 1L << 23;  // points per internal pass (bounds the scratch: 2.6 GB)

struct ParamCov {
  int numParams;
  uint32_t nonzero[kMaxParams * kMaxParams / 32];  // bit k: c[k] is not zero (a NaN entry counts as set)
  double c[kMaxParams * kMaxParams];  // getParameterCovariance(i, j) = c[i * numParams + j]
};

constexpr uint32_t kErrBits = USGSCSM_B200_ST_ERROR_ALGORITHM | USGSCSM_B200_ST_ERROR_UNKNOWN |
                              USGSCSM_B200_ST_ERROR_OTHER;

// determineSensorCovarianceInImageSpace LS:1025-1054 (PF:875-903, SAR:1897-1923)
__device__ inline void sensorCovariance(const ParamCov &C, const double *__restrict__ sp,
                                        double s[4]) {
  s[0] = s[1] = s[2] = s[3] = 0.0;
  const int P = C.numParams;
  for (int i = 0; i < P; i++) {
    const double li = sp[2 * i], si = sp[2 * i + 1];
    for (int j = 0; j < P; j++) {
      const double cij = C.c[i * P + j];
      const double lj = sp[2 * j], sj = sp[2 * j + 1];
      s[0] += li * cij * lj;
      s[1] += si * cij * lj;
      s[2] += li * cij * sj;
      s[3] += si * cij * sj;
    }
  }
}

// The same sums for a parameter count known at compile time: the point's 2 P partials are loaded
// once (128-bit loads) and stay in registers, the covariance entries are constant-bank operands,
// and the terms of entries that are exactly zero are skipped when every partial of the point is
// finite -- such a term is li * 0 * lj = +-0, and adding +-0 leaves a sum that started at +0
// unchanged, so the result has the same bits (with a non-finite partial the term is NaN and
// nothing is skipped).  A model straight from an ISD has a diagonal covariance: 16 terms of 256.
template <int P>
__device__ inline void sensorCovarianceFixed(const ParamCov &C, const double *__restrict__ spPoint,
                                             double s[4]) {
  double sp[2 * P];
  const double2 *q = reinterpret_cast<const double2 *>(spPoint);
  bool finite = true;
#pragma unroll
  for (int k = 0; k < P; k++) {
    const double2 v = __ldg(q + k);
    sp[2 * k] = v.x;
    sp[2 * k + 1] = v.y;
    finite = finite && isfinite(v.x) && isfinite(v.y);
  }
  s[0] = s[1] = s[2] = s[3] = 0.0;
#pragma unroll
  for (int i = 0; i < P; i++) {
#pragma unroll
    for (int j = 0; j < P; j++) {
      const int k = i * P + j;
      if (finite && !((C.nonzero[k >> 5] >> (k & 31)) & 1u)) continue;
      const double cij = C.c[k];
      s[0] += sp[2 * i] * cij * sp[2 * j];
      s[1] += sp[2 * i + 1] * cij * sp[2 * j];
      s[2] += sp[2 * i] * cij * sp[2 * j + 1];
      s[3] += sp[2 * i + 1] * cij * sp[2 * j + 1];
    }
  }
}

__device__ inline void sensorCovarianceOf(const ParamCov &C, const double *__restrict__ spAll, long i,
                                          double s[4]) {
  const double *sp = spAll + i * 2 * C.numParams;
  if (C.numParams == 16)
    sensorCovarianceFixed<16>(C, sp, s);
  else if (C.numParams == 6)
    sensorCovarianceFixed<6>(C, sp, s);
  else
    sensorCovariance(C, sp, s);
}

// LS:930-961
__global__ void groundToImageCovarKernel(long n, const __grid_constant__ ParamCov C,
                                         const double *__restrict__ groundCov,
                                         const double *__restrict__ prtAll,
                                         const double *__restrict__ spAll,
                                         const uint32_t *__restrict__ stPartials,
                                         const uint32_t *__restrict__ stSensor,
                                         double *__restrict__ line, double *__restrict__ samp,
                                         double *__restrict__ imageCov,
                                         double *__restrict__ achieved,
                                         uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double *prt = prtAll + i * 6;
    const double *cov = groundCov + i * 9;
    const double ltx = prt[0] * cov[0] + prt[1] * cov[3] + prt[2] * cov[6];
    const double lty = prt[0] * cov[1] + prt[1] * cov[4] + prt[2] * cov[7];
    const double ltz = prt[0] * cov[2] + prt[1] * cov[5] + prt[2] * cov[8];
    const double stx = prt[3] * cov[0] + prt[4] * cov[3] + prt[5] * cov[6];
    const double sty = prt[3] * cov[1] + prt[4] * cov[4] + prt[5] * cov[7];
    const double stz = prt[3] * cov[2] + prt[4] * cov[5] + prt[5] * cov[8];
    double g[4];
    g[0] = ltx * prt[0] + lty * prt[1] + ltz * prt[2];
    g[1] = ltx * prt[3] + lty * prt[4] + ltz * prt[5];
    g[2] = stx * prt[0] + sty * prt[1] + stz * prt[2];
    g[3] = stx * prt[3] + sty * prt[4] + stz * prt[5];
    double s[4];
    sensorCovarianceOf(C, spAll, i, s);
    uint32_t st = status[i];
    const uint32_t err = (st | stPartials[i] | stSensor[i]) & kErrBits;
    bool finite = true;
    for (int k = 0; k < 4; k++) {
      double v = g[k] + 0.0 + s[k];  // + getUnmodeledError (zero for these models)
      if (err) v = nan("");
      finite = finite && isfinite(v);
      imageCov[i * 4 + k] = v;
    }
    if (err) {  // one of the scalar calls inside would have thrown: nothing is returned
      st = err | USGSCSM_B200_ST_OUT_OF_IMAGE;  // NaN is not inside the image (derived bit)
      line[i] = samp[i] = nan("");
      if (achieved) achieved[i] = nan("");
    }
    if (!finite) st |= USGSCSM_B200_ST_NONFINITE;
    status[i] = st;
  }
}

// the three displaced inputs of LS:1092-1113
__global__ void displacedInputsKernel(long n, const double *__restrict__ line,
                                      const double *__restrict__ samp,
                                      const double *__restrict__ height, int heightStride,
                                      double deltaImage, double deltaGround,
                                      double *__restrict__ line1, double *__restrict__ samp1,
                                      double *__restrict__ height1) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    line1[i] = line[i] + deltaImage;
    samp1[i] = samp[i] + deltaImage;
    height1[i] = height[i * heightStride] + deltaGround;
  }
}

// LS:1094-1154
__global__ void imageToGroundCovarKernel(long n, const __grid_constant__ ParamCov C,
                                         const double *__restrict__ imageCovIn,
                                         const double *__restrict__ heightVar, int hvStride,
                                         double deltaImage, double deltaGround,
                                         const double *__restrict__ gl,   // 3 x n: x, y, z
                                         const double *__restrict__ gs,
                                         const double *__restrict__ gh,
                                         const double *__restrict__ spAll,
                                         const uint32_t *__restrict__ stOthers,  // 4 x n
                                         long ld, double *__restrict__ x, double *__restrict__ y,
                                         double *__restrict__ z, double *__restrict__ groundCov,
                                         double *__restrict__ achieved,
                                         uint32_t *__restrict__ status) {
  const long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double gx = x[i], gy = y[i], gz = z[i];
    const double xpl = (gl[i] - gx) / deltaImage, ypl = (gl[ld + i] - gy) / deltaImage,
                 zpl = (gl[2 * ld + i] - gz) / deltaImage;
    const double xps = (gs[i] - gx) / deltaImage, yps = (gs[ld + i] - gy) / deltaImage,
                 zps = (gs[2 * ld + i] - gz) / deltaImage;
    const double xph = (gh[i] - gx) / deltaGround, yph = (gh[ld + i] - gy) / deltaGround,
                 zph = (gh[2 * ld + i] - gz) / deltaGround;
    double s[4], ic[4];
    sensorCovarianceOf(C, spAll, i, s);
    for (int k = 0; k < 4; k++) ic[k] = imageCovIn[i * 4 + k] + s[k] + 0.0;
    const double hv = heightVar[i * hvStride];
    const double t00 = xpl * ic[0] + xps * ic[2], t01 = xpl * ic[1] + xps * ic[3], t02 = xph * hv;
    const double t10 = ypl * ic[0] + yps * ic[2], t11 = ypl * ic[1] + yps * ic[3], t12 = yph * hv;
    const double t20 = zpl * ic[0] + zps * ic[2], t21 = zpl * ic[1] + zps * ic[3], t22 = zph * hv;
    double g[9];
    g[0] = t00 * xpl + t01 * xps + t02 * xph;
    g[1] = t00 * ypl + t01 * yps + t02 * yph;
    g[2] = t00 * zpl + t01 * zps + t02 * zph;
    g[3] = t10 * xpl + t11 * xps + t12 * xph;
    g[4] = t10 * ypl + t11 * yps + t12 * yph;
    g[5] = t10 * zpl + t11 * zps + t12 * zph;
    g[6] = t20 * xpl + t21 * xps + t22 * xph;
    g[7] = t20 * ypl + t21 * yps + t22 * yph;
    g[8] = t20 * zpl + t21 * zps + t22 * zph;
    uint32_t st = status[i];
    const uint32_t err =
        (st | stOthers[i] | stOthers[ld + i] | stOthers[2 * ld + i] | stOthers[3 * ld + i]) & kErrBits;
    bool finite = true;
    for (int k = 0; k < 9; k++) {
      double v = err ? nan("") : g[k];
      finite = finite && isfinite(v);
      groundCov[i * 9 + k] = v;
    }
    if (err) {
      st = err;
      x[i] = y[i] = z[i] = nan("");
      if (achieved) achieved[i] = nan("");
    }
    if (!finite) st |= USGSCSM_B200_ST_NONFINITE;
    status[i] = st;
  }
}

ParamCov paramCov(const usgscsm_b200_model *m) {
  ParamCov C{};
  C.numParams = (int)m->params.size();
  if (C.numParams > kMaxParams ||
      m->covariance.size() != (size_t)C.numParams * (size_t)C.numParams)
    throw std::invalid_argument("m_covariance does not hold numParameters x numParameters entries");
  for (size_t k = 0; k < m->covariance.size(); k++) {
    C.c[k] = m->covariance[k];
    if (!(C.c[k] == 0.0)) C.nonzero[k >> 5] |= 1u << (k & 31);  // NaN: set
  }
  return C;
}

int gridFor(long n) { return (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8); }

struct StreamScratch {  // stream-ordered scratch released on scope exit
  cudaStream_t stream;
  std::vector<void *> ptrs;
  explicit StreamScratch(cudaStream_t s) : stream(s) {}
  ~StreamScratch() {
    for (void *p : ptrs) cudaFreeAsync(p, stream);
  }
  template <class T>
  T *get(size_t count) {
    void *p = nullptr;
    UB_CUDA(cudaMallocAsync(&p, (count ? count : 1) * sizeof(T), stream));
    ptrs.push_back(p);
    return static_cast<T *>(p);
  }
};

}  // namespace

int covarCheck(const usgscsm_b200_model *m, int64_t n) {
  if (!m || n < 0) return fail(USGSCSM_B200_EINVAL, "bad argument");
  if (m->kind == USGSCSM_B200_MODEL_FRAME)
    return fail(USGSCSM_B200_EUNSUPPORTED,
                "the frame model has no covariance propagation (stubs in the reference, "
                "UsgsAstroFrameSensorModel.cpp:234-278, 374-395)");
  return USGSCSM_B200_OK;
}

int groundToImageCovarDev(usgscsm_b200_model *m, long n, const double *x, const double *y,
                          const double *z, const double *groundCov, double prec, double *line,
                          double *samp, double *imageCov, double *achieved, uint32_t *status,
                          cudaStream_t stream) {
  const ParamCov C = paramCov(m);
  const int P = C.numParams;
  const long B = n < kCovarBatch ? n : kCovarBatch;
  StreamScratch s(stream);
  double *prt = s.get<double>((size_t)B * 6), *sp = s.get<double>((size_t)B * 2 * P);
  uint32_t *stP = s.get<uint32_t>((size_t)B), *stS = s.get<uint32_t>((size_t)B);
  uint32_t *stOwn = status ? nullptr : s.get<uint32_t>((size_t)B);
  for (long off = 0; off < n; off += B) {
    const long cnt = n - off < B ? n - off : B;
    uint32_t *st = status ? status + off : stOwn;
    double *ach = achieved ? achieved + off : nullptr;
    UB_CUDA(launchGroundToImage(*m, cnt, x + off, y + off, z + off, prec, line + off, samp + off,
                                ach, st, stream));
    UB_CUDA(launchGroundPartials(*m, cnt, x + off, y + off, z + off, prt, stP, stream));
    // computeAllSensorPartials(gp): default arguments, i.e. precision 0.001 (LS:1034-1035): it
    // starts with groundToImage(gp, 0.001) -- the search just made when the caller's precision is
    // 0.001 too (the default), whose image point is then handed over instead of searched again
    // (its status is in `st`, which the combining kernel reads beside stS)
    const bool sameBase = prec == 0.001;
    const int rcPartials =
        allSensorPartialsDev(m, P, nullptr, cnt, x + off, y + off, z + off, sameBase ? line + off : nullptr,
                             sameBase ? samp + off : nullptr, 0.001, sp, nullptr, stS, stream);
    if (rcPartials != USGSCSM_B200_OK) return rcPartials;
    groundToImageCovarKernel<<<gridFor(cnt), 256, 0, stream>>>(
        cnt, C, groundCov + off * 9, prt, sp, stP, stS, line + off, samp + off, imageCov + off * 4,
        ach, st);
    countLaunch();
    UB_CUDA(cudaGetLastError());
  }
  return USGSCSM_B200_OK;
}

int imageToGroundCovarDev(usgscsm_b200_model *m, long n, const double *line, const double *samp,
                          const double *imageCov, const double *height, int heightStride,
                          const double *heightVar, int hvStride, double prec, double *x, double *y,
                          double *z, double *groundCov, double *achieved, uint32_t *status,
                          cudaStream_t stream) {
  const ParamCov C = paramCov(m);
  const int P = C.numParams;
  const double deltaImage = 1.0;  // DELTA_IMAGE LS:1086
  // DELTA_GROUND: m_gsd (LS:1087, PF:920), m_scaledPixelWidth (SAR:873)
  const double deltaGround =
      m->kind == USGSCSM_B200_MODEL_SAR ? m->sar.scaledPixelWidth : m->ls.gsd;
  const long B = n < kCovarBatch ? n : kCovarBatch;
  const int hs = heightStride ? 1 : 0, hvs = hvStride ? 1 : 0;
  StreamScratch s(stream);
  double *l1 = s.get<double>((size_t)B), *s1 = s.get<double>((size_t)B),
         *h1 = s.get<double>((size_t)B);
  double *gl = s.get<double>((size_t)B * 3), *gs = s.get<double>((size_t)B * 3),
         *gh = s.get<double>((size_t)B * 3), *sp = s.get<double>((size_t)B * 2 * P);
  uint32_t *stO = s.get<uint32_t>((size_t)B * 4);
  uint32_t *stOwn = status ? nullptr : s.get<uint32_t>((size_t)B);
  for (long off = 0; off < n; off += B) {
    const long cnt = n - off < B ? n - off : B;
    uint32_t *st = status ? status + off : stOwn;
    double *ach = achieved ? achieved + off : nullptr;
    const double *h = height + off * hs;
    UB_CUDA(launchImageToGround(*m, cnt, line + off, samp + off, nullptr, h, hs, prec, x + off,
                                y + off, z + off, ach, st, stream));
    displacedInputsKernel<<<gridFor(cnt), 256, 0, stream>>>(cnt, line + off, samp + off, h, hs,
                                                            deltaImage, deltaGround, l1, s1, h1);
    countLaunch();
    UB_CUDA(cudaGetLastError());
    UB_CUDA(launchImageToGround(*m, cnt, l1, samp + off, nullptr, h, hs, prec, gl, gl + B, gl + 2 * B,
                                nullptr, stO, stream));
    UB_CUDA(launchImageToGround(*m, cnt, line + off, s1, nullptr, h, hs, prec, gs, gs + B, gs + 2 * B,
                                nullptr, stO + B, stream));
    UB_CUDA(launchImageToGround(*m, cnt, line + off, samp + off, nullptr, h1, 1, prec, gh, gh + B,
                                gh + 2 * B, nullptr, stO + 2 * B, stream));
    allSensorPartialsDev(m, P, nullptr, cnt, x + off, y + off, z + off, nullptr, nullptr, 0.001, sp,
                         nullptr, stO + 3 * B, stream);
    imageToGroundCovarKernel<<<gridFor(cnt), 256, 0, stream>>>(
        cnt, C, imageCov + off * 4, heightVar + off * hvs, hvs, deltaImage, deltaGround, gl, gs, gh,
        sp, stO, B, x + off, y + off, z + off, groundCov + off * 9, ach, st);
    countLaunch();
    UB_CUDA(cudaGetLastError());
  }
  return USGSCSM_B200_OK;
}

}  // namespace ub200

using namespace ub200;

extern "C" {

int usgscsm_b200_ground_to_image_covar_dev(usgscsm_b200_model *m, int64_t n, const double *x,
                                           const double *y, const double *z,
                                           const double *ground_cov, double prec, double *line,
                                           double *samp, double *image_cov, double *achieved,
                                           uint32_t *status, void *stream) {
  int c = covarCheck(m, n);
  if (c != USGSCSM_B200_OK || n == 0) return c;
  if (!x || !y || !z || !ground_cov || !line || !samp || !image_cov)
    return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    return groundToImageCovarDev(m, (long)n, x, y, z, ground_cov, prec, line, samp, image_cov,
                                 achieved, status, static_cast<cudaStream_t>(stream));
  });
}

int usgscsm_b200_image_to_ground_covar_dev(usgscsm_b200_model *m, int64_t n, const double *line,
                                           const double *samp, const double *image_cov,
                                           const double *height, int height_stride,
                                           const double *height_variance, int variance_stride,
                                           double prec, double *x, double *y, double *z,
                                           double *ground_cov, double *achieved, uint32_t *status,
                                           void *stream) {
  int c = covarCheck(m, n);
  if (c != USGSCSM_B200_OK || n == 0) return c;
  if (!line || !samp || !image_cov || !height || !height_variance || !x || !y || !z || !ground_cov)
    return fail(USGSCSM_B200_EINVAL, "null array");
  return guarded([&]() -> int {
    UB_CUDA(cudaSetDevice(m->device));
    return imageToGroundCovarDev(m, (long)n, line, samp, image_cov, height, height_stride,
                                 height_variance, variance_stride, prec, x, y, z, ground_cov,
                                 achieved, status, static_cast<cudaStream_t>(stream));
  });
}

// getParameterCovariance / setParameterCovariance (LS:1468-1497, PF:1207-1230, SAR:1604-1623,
// FR:1289-1320)
int usgscsm_b200_model_get_parameter_covariance(const usgscsm_b200_model *m, int index1,
                                                int index2, double *value) {
  if (!m || !value) return fail(USGSCSM_B200_EINVAL, "null argument");
  const int P = (int)m->params.size();
  if (index1 < 0 || index2 < 0 || index1 >= P || index2 >= P ||
      m->covariance.size() != (size_t)P * (size_t)P)
    return fail(USGSCSM_B200_EINVAL, "parameter index out of range");
  *value = m->covariance[(size_t)index1 * P + index2];
  return USGSCSM_B200_OK;
}

int usgscsm_b200_model_set_parameter_covariance(usgscsm_b200_model *m, int index1, int index2,
                                                double value) {
  if (!m) return fail(USGSCSM_B200_EINVAL, "null argument");
  const int P = (int)m->params.size();
  if (index1 < 0 || index2 < 0 || index1 >= P || index2 >= P)
    return fail(USGSCSM_B200_EINVAL, "parameter index out of range");
  if (m->covariance.size() != (size_t)P * (size_t)P) m->covariance.assign((size_t)P * P, 0.0);
  m->covariance[(size_t)index1 * P + index2] = value;
  return USGSCSM_B200_OK;
}

}  // extern "C"
