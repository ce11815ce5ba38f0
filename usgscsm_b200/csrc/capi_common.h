// capi_common.h -- plumbing shared by the translation units that implement the C-ABI
// (capi.cu: models, device-pointer entry points, building blocks; capi_host.cu: host-pointer
// entry points with their copy pipeline).
#ifndef USGSCSM_B200_CAPI_COMMON_H_
#define USGSCSM_B200_CAPI_COMMON_H_

#include <stdexcept>
#include <string>

#include "../../include/usgscsm_b200.h"
#include "model.h"

namespace ub200 {

// Message of the last failure on the calling thread (usgscsm_b200_last_error).
std::string &lastError();

// Run f(); map what it throws to a negative USGSCSM_B200_E* code and remember the message.
template <class F>
int guarded(F &&f) {
  try {
    return f();
  } catch (const CudaError &e) {
    lastError() = e.what();
    return USGSCSM_B200_ECUDA;
  } catch (const StateError &e) {
    lastError() = e.what();
    return USGSCSM_B200_ESTATE;
  } catch (const std::invalid_argument &e) {
    lastError() = e.what();
    return USGSCSM_B200_EINVAL;
  } catch (const std::exception &e) {
    lastError() = e.what();
    return USGSCSM_B200_ESTATE;
  }
}

inline int fail(int code, const char *msg) {
  lastError() = msg;
  return code;
}

int deviceCount();

// Dispatch on the model kind; device pointers, enqueue on `stream` (capi.cu).
cudaError_t launchGroundToImage(Model &m, long n, const double *x, const double *y,
                                const double *z, double prec, double *line, double *samp,
                                double *achieved, uint32_t *status, cudaStream_t stream);
cudaError_t launchImageToGround(Model &m, long n, const double *line, const double *samp,
                                const GridSpec *grid, const double *height, int heightStride,
                                double prec, double *x, double *y, double *z, double *achieved,
                                uint32_t *status, cudaStream_t stream);
cudaError_t launchRemoteLocus(Model &m, long n, const double *line, const double *samp,
                              double prec, const LocusOut &o, cudaStream_t stream);
cudaError_t launchProximateLocus(Model &m, long n, const double *line, const double *samp,
                                 const double *x, const double *y, const double *z, double prec,
                                 const LocusOut &o, cudaStream_t stream);
cudaError_t launchGroundPartials(Model &m, long n, const double *x, const double *y,
                                 const double *z, double *partials, uint32_t *status,
                                 cudaStream_t stream);
// computeSensorPartials over device arrays; throws CudaError.  img_line / img_samp nullable.
int sensorPartialsDev(usgscsm_b200_model *m, int index, long n, const double *x, const double *y,
                      const double *z, const double *img_line, const double *img_samp,
                      double prec, double *dline, double *dsamp, double *achieved,
                      uint32_t *status, cudaStream_t stream);
int sensorPartialsCheck(usgscsm_b200_model *m, int index, int64_t n, const void *x, const void *y,
                        const void *z, const void *dline, const void *dsamp);

int allSensorPartialsDev(usgscsm_b200_model *m, int n_indices, const int *indices, long n,
                         const double *x, const double *y, const double *z,
                         const double *img_line, const double *img_samp, double prec,
                         double *partials, double *achieved, uint32_t *status,
                         cudaStream_t stream);
int allSensorPartialsCheck(usgscsm_b200_model *m, int n_indices, const int *indices, int64_t n,
                           const void *x, const void *y, const void *z, const void *partials);

// illumination.cu
int illuminationDev(usgscsm_b200_model *m, long n, const double *x, const double *y,
                    const double *z, double *dx, double *dy, double *dz, uint32_t *status,
                    cudaStream_t stream);
// covar.cu
int covarCheck(const usgscsm_b200_model *m, int64_t n);
int groundToImageCovarDev(usgscsm_b200_model *m, long n, const double *x, const double *y,
                          const double *z, const double *groundCov, double prec, double *line,
                          double *samp, double *imageCov, double *achieved, uint32_t *status,
                          cudaStream_t stream);
int imageToGroundCovarDev(usgscsm_b200_model *m, long n, const double *line, const double *samp,
                          const double *imageCov, const double *height, int heightStride,
                          const double *heightVar, int hvStride, double prec, double *x, double *y,
                          double *z, double *groundCov, double *achieved, uint32_t *status,
                          cudaStream_t stream);

}  // namespace ub200
#endif
