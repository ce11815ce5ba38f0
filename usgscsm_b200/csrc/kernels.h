// kernels.h -- host-side declarations of the kernel launchers (one .cu per sensor model).
#ifndef USGSCSM_B200_KERNELS_H_
#define USGSCSM_B200_KERNELS_H_

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/usgscsm_b200.h"
#include "sensor_state.h"

namespace ub200 {

// Tables up to this size are staged into shared memory by every block (keeps >= 4 blocks
// of 256 threads resident per SM: 4 x 48 KB < 227 KB); larger tables (e.g. the 1.28 MB
// HiRISE quaternion table) are read through L1/L2, where neighbouring threads -- which
// need neighbouring times -- hit the same lines.
constexpr size_t kMaxStagedTableBytes = 48 * 1024;
// Persistent grid: resident blocks per SM the launchers size the grid for.
constexpr int kBlocksPerSm = 4;

struct LaunchShape {
  int blocks;
  int threads;
};

// Dense pixel grid: point k = r * nSamps + c  ->  (line0 + (row0 + r) * dline, samp0 + c * dsamp).
// row0: first grid row of this launch (a chunk of a larger grid), so that the line of a row is
// the same double whichever chunk it falls in.
struct GridSpec {
  double line0, samp0, dline, dsamp;
  long nSamps;
  long row0;
};

void countLaunch();

// Outputs of imageToRemoteImagingLocus (csm::EcefLocus = point + direction, csm/csm.h:420-445),
// SoA; achieved / status nullable.
struct LocusOut {
  double *px, *py, *pz, *dx, *dy, *dz, *achieved;
  uint32_t *status;
};
#ifdef __CUDACC__
__device__ __forceinline__ void storeLocus(const LocusOut &o, long i, const double p[3],
                                           const double d[3], double achieved, uint32_t st) {
  __stcs(o.px + i, p[0]);
  __stcs(o.py + i, p[1]);
  __stcs(o.pz + i, p[2]);
  __stcs(o.dx + i, d[0]);
  __stcs(o.dy + i, d[1]);
  __stcs(o.dz + i, d[2]);
  if (o.achieved) __stcs(o.achieved + i, achieved);
  if (o.status) __stcs(o.status + i, st);
}
// computeGroundPartials: the scalar call has no warning list and lets a thrown csm::Error
// escape, so only the error bits of its four groundToImage calls are observable; a point
// that would have thrown gets NaN partials.  partials: 6 doubles per point,
// [dLine/dX, dLine/dY, dLine/dZ, dSamp/dX, dSamp/dY, dSamp/dZ] (LS:1296-1298).
__device__ __forceinline__ void storePartials(double *__restrict__ partials,
                                              uint32_t *__restrict__ status, long i, double p[6],
                                              uint32_t st) {
  st &= (USGSCSM_B200_ST_ERROR_ALGORITHM | USGSCSM_B200_ST_ERROR_UNKNOWN |
         USGSCSM_B200_ST_ERROR_OTHER);
  bool finite = true;
#pragma unroll
  for (int k = 0; k < 6; k++) {
    if (st) p[k] = nan("");
    finite = finite && isfinite(p[k]);
    __stcs(partials + 6 * i + k, p[k]);
  }
  if (!finite) st |= USGSCSM_B200_ST_NONFINITE;
  if (status) __stcs(status + i, st);
}
#endif

// getImageTime / getSensorPosition / getSensorVelocity per image point; any array may be null.
struct SensorStateOut {
  double *time, *px, *py, *pz, *vx, *vy, *vz;
  uint32_t *status;
  int atTime;  // the input array holds image TIMES (getSensorPosition(double time)), not lines
};
#ifdef __CUDACC__
__device__ __forceinline__ void storeSensorState(const SensorStateOut &o, long i, double time,
                                                 const double pos[3], const double vel[3],
                                                 uint32_t st) {
  if (!(isfinite(time) && isfinite(pos[0]) && isfinite(pos[1]) && isfinite(pos[2]) &&
        isfinite(vel[0]) && isfinite(vel[1]) && isfinite(vel[2])))
    st |= USGSCSM_B200_ST_NONFINITE;
  if (o.time) o.time[i] = time;
  if (o.px) o.px[i] = pos[0];
  if (o.py) o.py[i] = pos[1];
  if (o.pz) o.pz[i] = pos[2];
  if (o.vx) o.vx[i] = vel[0];
  if (o.vy) o.vy[i] = vel[1];
  if (o.vz) o.vz[i] = vel[2];
  if (o.status) o.status[i] = st;
}
#endif
cudaError_t launchLsSensorState(const LsState &S, long n, const double *line,
                                const SensorStateOut &o, cudaStream_t stream);
cudaError_t launchPfSensorState(const LsState &S, long n, const double *line,
                                const SensorStateOut &o, cudaStream_t stream);
cudaError_t launchFrSensorState(const FrameState &S, long n, const double *line,
                                const double *samp, const SensorStateOut &o, cudaStream_t stream);
cudaError_t launchSarSensorState(const SarState &S, long n, const double *line,
                                 const SensorStateOut &o, cudaStream_t stream);

cudaError_t launchGroundPartialsLs(const LsState &S, int smCount, long n, const double *x,
                                   const double *y, const double *z, double *partials,
                                   uint32_t *status, cudaStream_t stream);
cudaError_t launchGroundPartialsPf(const LsState &S, int smCount, long n, const double *x,
                                   const double *y, const double *z, double *partials,
                                   uint32_t *status, cudaStream_t stream);
cudaError_t launchGroundPartialsFr(const FrameState &S, int smCount, long n, const double *x,
                                   const double *y, const double *z, double *partials,
                                   uint32_t *status, cudaStream_t stream);
cudaError_t launchGroundPartialsSar(const SarState &S, int smCount, long n, const double *x,
                                    const double *y, const double *z, double *partials,
                                    uint32_t *status, cudaStream_t stream);

cudaError_t launchLsGroundToImage(const LsState &S, int smCount, long n, const double *x,
                                  const double *y, const double *z, double prec, double *line,
                                  double *samp, double *achieved, uint32_t *status,
                                  cudaStream_t stream);
cudaError_t launchLsSensorPartialStep(const LsState &S, int smCount, long n, const double *x,
                                      const double *y, const double *z, double prec, const double *l0,
                                      const double *s0, const uint32_t *st0, double delta,
                                      double *partials, int numParams, int k, double *achieved,
                                      uint32_t *status, cudaStream_t stream);
cudaError_t launchLsImageToGround(const LsState &S, int smCount, long n, const double *line,
                                  const double *samp, const GridSpec *grid, const double *height,
                                  int heightStride, double prec, double *x, double *y, double *z,
                                  double *achieved, uint32_t *status, cudaStream_t stream);

// Dense-grid imageToGround with its row / column tables in a file-scope device array instead of
// stream-ordered scratch: no allocation of any kind (usgscsm_b200_image_to_ground_grid_static_dev).
// 2 Mi doubles = 16 MiB per GPU; a grid fits when rows * 21 + columns * 4 <= that.
constexpr long kStaticGridScratchDoubles = 1L << 21;
bool staticGridScratchFits(long nRows, long nCols);
cudaError_t launchLsImageToGroundGridStatic(const LsState &S, int smCount, const GridSpec &grid, long nRows,
                                            const double *height, int heightStride, double prec,
                                            double *x, double *y, double *z, double *achieved,
                                            uint32_t *status, cudaStream_t stream);

// computeAllSensorPartials of the line-scan model in one kernel (ls_all_partials.cu); false: not
// applicable, run the per-parameter launches
bool launchLsAllSensorPartials(const LsState &S, int smCount, int nIndices, const int *indices, long n,
                               const double *x, const double *y, const double *z,
                               const double *imgLine, const double *imgSamp, double prec,
                               double *partials, double *achieved, uint32_t *status,
                               cudaStream_t stream, cudaError_t &err);

cudaError_t launchLsRemoteLocus(const LsState &S, int smCount, long n, const double *line,
                                const double *samp, const LocusOut &o, cudaStream_t stream);
cudaError_t launchPfRemoteLocus(const LsState &S, int smCount, long n, const double *line,
                                const double *samp, const LocusOut &o, cudaStream_t stream);
cudaError_t launchFrRemoteLocus(const FrameState &S, int smCount, long n, const double *line,
                                const double *samp, const LocusOut &o, cudaStream_t stream);
cudaError_t launchSarRemoteLocus(const SarState &S, int smCount, long n, const double *line,
                                 const double *samp, double prec, const LocusOut &o,
                                 cudaStream_t stream);

cudaError_t launchLsProximateLocus(const LsState &S, int smCount, long n, const double *line,
                                   const double *samp, const double *x, const double *y,
                                   const double *z, double prec, const LocusOut &o,
                                   cudaStream_t stream);
cudaError_t launchPfProximateLocus(const LsState &S, int smCount, long n, const double *line,
                                   const double *samp, const double *x, const double *y,
                                   const double *z, double prec, const LocusOut &o,
                                   cudaStream_t stream);
cudaError_t launchFrProximateLocus(const FrameState &S, int smCount, long n, const double *line,
                                   const double *samp, const double *x, const double *y,
                                   const double *z, const LocusOut &o, cudaStream_t stream);
cudaError_t launchSarProximateLocus(const SarState &S, int smCount, long n, const double *line,
                                    const double *samp, const double *x, const double *y,
                                    const double *z, const LocusOut &o, cudaStream_t stream);

// push-frame (pf_kernels.cu)
cudaError_t launchPfFrameletOrientations(const LsState &S, double *out, cudaStream_t stream);
cudaError_t launchPfSensorPosition(const LsState &S, double time, double *xyz,
                                   cudaStream_t stream);
cudaError_t launchPfGroundToImage(const LsState &S, int smCount, long n, const double *x,
                                  const double *y, const double *z, double prec, double *line,
                                  double *samp, double *achieved, uint32_t *status,
                                  cudaStream_t stream);
cudaError_t launchPfImageToGround(const LsState &S, int smCount, long n, const double *line,
                                  const double *samp, const GridSpec *grid, const double *height,
                                  int heightStride, double prec, double *x, double *y, double *z,
                                  double *achieved, uint32_t *status, cudaStream_t stream);

// frame (fr_kernels.cu)
cudaError_t launchFrGroundToImage(const FrameState &S, int smCount, long n, const double *x,
                                  const double *y, const double *z, double prec, double *line,
                                  double *samp, double *achieved, uint32_t *status,
                                  cudaStream_t stream);
cudaError_t launchFrImageToGround(const FrameState &S, int smCount, long n, const double *line,
                                  const double *samp, const GridSpec *grid, const double *height,
                                  int heightStride, double prec, double *x, double *y, double *z,
                                  double *achieved, uint32_t *status, cudaStream_t stream);

// SAR (sar_kernels.cu)
cudaError_t launchSarGroundToImage(const SarState &S, int smCount, long n, const double *x,
                                   const double *y, const double *z, double prec, double *line,
                                   double *samp, double *achieved, uint32_t *status,
                                   cudaStream_t stream);
cudaError_t launchSarImageToGround(const SarState &S, int smCount, long n, const double *line,
                                   const double *samp, const GridSpec *grid, const double *height,
                                   int heightStride, double prec, double *x, double *y, double *z,
                                   double *achieved, uint32_t *status, cudaStream_t stream);

// building blocks (blocks_kernels.cu)
cudaError_t launchDistortion(bool apply, const FocalPlane &fp, double precision, long n,
                             const double *inX, const double *inY, double *outX, double *outY,
                             uint32_t *status, cudaStream_t stream);
cudaError_t launchLagrange(const UniformTable &tab, const LagrangeDivisors &lag, int vlen, long n,
                           const double *times, double *out, cudaStream_t stream);
cudaError_t launchFocalPlane(const FocalPlane &fp, int toPixel, long n, const double *a, const double *b,
                             double *outA, double *outB, cudaStream_t stream);
cudaError_t launchLookVector(double zDirection, double focalLength, long n, const double *ux,
                             const double *uy, double *look, cudaStream_t stream);
cudaError_t launchCubic(const double coeffs[4], int mode, double eps, long n, const double *a,
                        const double *b, double *out0, double *out1, uint32_t *status,
                        cudaStream_t stream);
cudaError_t launchJitter(const FrameState &S, int remove, double tolerance, int maxIts, long n,
                         const double *line, const double *sample, double *outLine,
                         double *outSample, cudaStream_t stream);
cudaError_t launchAosSplit(int ncomp, long n, const double *aos, double *c0, double *c1,
                           double *c2, cudaStream_t stream);
cudaError_t launchAosJoin(int ncomp, long n, const double *c0, const double *c1, const double *c2,
                          double *aos, cudaStream_t stream);
cudaError_t launchSensorPartialsCombine(long n, const double *l0, const double *s0,
                                        const uint32_t *st0, const double *l1, const double *s1,
                                        double delta, double *dline, double *dsamp,
                                        double *achieved, uint32_t *status, cudaStream_t stream);
cudaError_t launchSensorPartialsStep(long n, const double *l0, const double *s0,
                                     const uint32_t *st0, const double *l1, const double *s1,
                                     const double *achP, const uint32_t *stP, double delta,
                                     double *partials, int numParams, int k, double *achieved,
                                     uint32_t *status, cudaStream_t stream);
cudaError_t launchReciprocals(int n, const double *b, double *r, cudaStream_t stream);
cudaError_t launchDivisionSelfTest(unsigned long long seed, int blocks, long perThread,
                                   unsigned long long *mismatches, cudaStream_t stream);

cudaError_t launchLsSensorPosition(const LsState &S, int n, const double *line, double *xyz,
                                   cudaStream_t stream);
cudaError_t launchEllipsoidElevation(const Ellipsoid &ell, int n, const double *xyz, double prec,
                                     double *h, cudaStream_t stream);

}  // namespace ub200
#endif
