// geogrid.h -- georeferenced raster grids around the hot path.
//
// What orthorectification callers do on either side of groundToImage / imageToGround: a DEM
// (or map-projected image) pixel -> map coordinates by the raster's geotransform
// (pixelToMeter / meterToPixel, src/Utilities.cpp:2720-2745) -> geographic coordinates by the
// inverse map projection -> body-fixed cartesian coordinates (the chain
// UsgsAstroProjectedSensorModel::imageToGround runs through PROJ,
// src/UsgsAstroProjectedSensorModel.cpp:329-365, 128-147), and back
// (::groundToImage, :289-321).  The projection arithmetic is PROJ's (vendored with the reference
// under PROJ/src): `cart` conversion PROJ/src/conversions/cart.cpp:107-199, `eqc`
// PROJ/src/projections/eqc.cpp, `sinu` PROJ/src/projections/gn_sinu.cpp:24-98 with the meridional
// distance series of PROJ/src/mlfn.cpp, generic scaling / offsets PROJ/src/fwd.cpp, inv.cpp,
// longitude wrapping PROJ/src/adjlon.cpp.  Three projections are built, the ones planetary DEMs
// and the reference's own projected fixtures use: longlat (degrees), eqc, sinu.
#ifndef USGSCSM_B200_GEOGRID_H_
#define USGSCSM_B200_GEOGRID_H_

#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/usgscsm_b200.h"

namespace ub200 {

constexpr int kMlfnCoeffs = 13;  // 2 * Lmax + 1 (PROJ/src/mlfn.cpp)

// Device-ready description of the raster's georeferencing.
struct GeoGrid {
  double gt[6];     // GDAL geotransform; gt[2], gt[4] are ignored as in the reference (U:2720-2745)
  int projection;   // USGSCSM_B200_PROJ_*
  int ellipsoidal;  // projection evaluated on the ellipsoid (sinu with es != 0)
  double a, b;      // ellipsoid of the geographic -> cartesian step
  double es;        // 1 - b^2 / a^2
  double oneEs;     // 1 - es
  double ra;        // 1 / a  (PROJ P->ra)
  double projA;     // the projection's own semi-major axis (P->a of the projected CRS)
  double projRa;
  double projEs;    // eqc forces 0 (eqc.cpp:89); sinu keeps it
  double x0, y0;    // false easting / northing, metres
  double lam0;      // central meridian, radians
  double phi0;      // latitude of origin, radians (eqc)
  double rc;        // cos(lat_ts) (eqc)
  double en[kMlfnCoeffs];  // pj_enfn(n) for ellipsoidal sinu
};

// Parses the subset of PROJ strings the engine evaluates itself:
//   +proj=longlat|latlong|lonlat|latlon, +proj=eqc [+lat_ts +lat_0 +lon_0], +proj=sinu [+lon_0],
//   +x_0 +y_0, +R | +a +b | +a +rf | +a +f, +units=m, +no_defs, +type=crs.
// Throws std::invalid_argument (-> USGSCSM_B200_EUNSUPPORTED at the C-ABI) for anything else.
// a, b of the cartesian step default to the projection's ellipsoid; the projected sensor model
// passes the CAMERA's radii instead (src/UsgsAstroProjectedSensorModel.cpp:131-139).
GeoGrid geoGridFromProjString(const std::string &proj, const double gt[6]);
GeoGrid geoGridFromC(const usgscsm_b200_geogrid &g);

// Separable projections (longlat, eqc): latitude depends on the raster row only and longitude on
// the column only, so the trigonometry of a whole window is 4 values per row and 2 per column.
// They are evaluated HERE, on the host, with the C library's sin / cos -- the functions PROJ
// calls -- so that the per-pixel kernel is left with + and * and reproduces cart.cpp:137-139 bit
// for bit.  rowTab: per row {N + 0, N * (1 - es), cos(phi), sin(phi)}; colTab: per column
// {cos(lam), sin(lam)}.  Row r of the window is raster line line0 + r * dline (pixel centres:
// line0 = first row + 0.5), column c is sample samp0 + c * dsamp.
bool geoGridSeparable(const GeoGrid &g);
void geoGridTables(const GeoGrid &g, double line0, double dline, int64_t nRows, double samp0,
                   double dsamp, int64_t nCols, std::vector<double> &rowTab,
                   std::vector<double> &colTab);

}  // namespace ub200
#endif
