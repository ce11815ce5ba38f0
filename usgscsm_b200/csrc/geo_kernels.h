// geo_kernels.h -- host-side declarations of the georeferenced-grid kernels (geo_kernels.cu) and
// the PODs they take.
#ifndef USGSCSM_B200_GEO_KERNELS_H_
#define USGSCSM_B200_GEO_KERNELS_H_

#include <cuda_runtime.h>
#include <stdint.h>

#include "geogrid.h"
#include "sensor_state.h"

namespace ub200 {

// DEM samples of one raster window as the kernels see them.
struct DemWindow {
  const void *samples;  // first sample of the window (device)
  int sampleType;       // USGSCSM_B200_DEM_*
  int hasNodata;
  double nodata;        // raw sample value
  double scale, offset; // height = raw * scale + offset
  long rowStride;       // samples between consecutive rows
  long nCols;           // window columns
  // separable projections: host-built tables of the window (geoGridTables); null otherwise
  const double *rowTab;  // 4 doubles per window row
  const double *colTab;  // 2 doubles per window column
  // per-pixel projection (sinu)
  GeoGrid grid;
  double line0, samp0;  // raster coordinates of the window's first pixel centre
};

// A whole DEM raster in device memory (bilinear sampling by the DEM-iterated intersection).
struct DemRaster {
  const void *samples;
  int sampleType;
  int hasNodata;
  double nodata;
  double scale, offset;
  long nRows, nCols;
  GeoGrid grid;
};

cudaError_t launchDemGroundPoints(const DemWindow &w, int smCount, long n, double *x, double *y,
                                  double *z, uint32_t *status, cudaStream_t stream);
cudaError_t launchDemNodataFixup(const DemWindow &w, int smCount, long n, double *line, double *samp,
                                 double *achieved, uint32_t *status, cudaStream_t stream);
cudaError_t launchGridImageToGround(const GeoGrid &g, int smCount, long n, const double *line,
                                    const double *samp, const double *height, int heightStride,
                                    double *x, double *y, double *z, uint32_t *status,
                                    cudaStream_t stream);
cudaError_t launchGridGroundToImage(const GeoGrid &g, int smCount, long n, const double *x,
                                    const double *y, const double *z, double *line, double *samp,
                                    double *height, uint32_t *status, cudaStream_t stream);
cudaError_t launchDemHeightUpdate(const DemRaster &d, int smCount, long n, const double *x,
                                  const double *y, const double *z, double *h,
                                  const uint32_t *i2gStatus, uint32_t *state, double tolerance,
                                  int last, cudaStream_t stream);
cudaError_t launchFill(int smCount, long n, double v, double *a, uint32_t *st, cudaStream_t stream);
// ls_kernels.cu: line-scan groundToImage with the ground points of a DEM window generated in
// the kernel
cudaError_t launchLsGroundToImageDem(const LsState &S, int smCount, long n, const DemWindow &w,
                                     double prec, double *line, double *samp, double *achieved,
                                     uint32_t *status, cudaStream_t stream);

}  // namespace ub200
#endif
