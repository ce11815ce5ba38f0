// geo_device.cuh -- device functions of the georeferenced-grid support (geogrid.h): raster pixel
// <-> map coordinates <-> geographic <-> body-fixed cartesian, and DEM sample access.
// Arithmetic follows PROJ as vendored with the reference (files named at each function); the
// translation units are compiled without FMA contraction like the rest of the engine.
#ifndef USGSCSM_B200_GEO_DEVICE_CUH_
#define USGSCSM_B200_GEO_DEVICE_CUH_

#include "device_math.cuh"
#include "geo_kernels.h"
#include "geogrid.h"

namespace ub200 {

constexpr double kGeoDegToRad = 0.017453292519943296;  // PROJ/src/proj.h DEG_TO_RAD
constexpr double kGeoRadToDeg = 57.295779513082321;    // RAD_TO_DEG
constexpr double kGeoPi = 3.14159265358979323846;
constexpr double kGeoHalfPi = 1.57079632679489661923;
constexpr double kGeoTwoPi = 6.28318530717958647693;

// PROJ/src/adjlon.cpp:7-23
UB_DEV double geoAdjlon(double longitude) {
  if (fabs(longitude) < kGeoPi + 1e-12) return longitude;
  longitude += kGeoPi;
  longitude -= kGeoTwoPi * floor(longitude / kGeoTwoPi);
  longitude -= kGeoPi;
  return longitude;
}

// clenshaw / pj_mlfn / pj_inv_mlfn, PROJ/src/mlfn.cpp:21-32, 75-83
UB_DEV double geoClenshaw(double szeta, double czeta, const double *c, int K) {
  double u0 = 0, u1 = 0, X = 2 * (czeta - szeta) * (czeta + szeta);
  for (; K > 0;) {
    double t = X * u0 - u1 + c[--K];
    u1 = u0;
    u0 = t;
  }
  return 2 * szeta * czeta * u0;
}
UB_DEV double geoMlfn(double phi, double sphi, double cphi, const double *en) {
  return en[0] * (phi + geoClenshaw(sphi, cphi, en + 1, 6));
}
UB_DEV double geoInvMlfn(double mu, const double *en) {
  mu /= en[0];
  return mu + geoClenshaw(sin(mu), cos(mu), en + 1 + 6, 6);
}

// Map coordinates (metres, or degrees for longlat) -> geographic radians: the inverse
// projection step of the PJ_FWD direction of m_isdProj2ecefProj
// (src/UsgsAstroProjectedSensorModel.cpp:347-350): inv_prepare PROJ/src/inv.cpp:77-97, the
// projection's inverse, inv_finalize :113-121.  False when outside the projection's domain.
UB_DEV bool geoProjInverse(const GeoGrid &g, double mx, double my, double &lam, double &phi) {
  if (g.projection == USGSCSM_B200_PROJ_LONGLAT) {
    lam = mx * kGeoDegToRad;
    phi = my * kGeoDegToRad;
    return true;
  }
  double x = (mx - g.x0) * g.projRa;
  double y = (my - g.y0) * g.projRa;
  bool ok = true;
  if (g.projection == USGSCSM_B200_PROJ_EQC) {  // eqc.cpp:52-61
    lam = x / g.rc;
    phi = y + g.phi0;
  } else if (!g.ellipsoidal) {  // gn_sinu.cpp s_inverse :87-97 with m = 0, n = 1, C_x = C_y = 1
    phi = y;
    lam = x / cos(y);
  } else {  // gn_sinu.cpp e_inverse :36-52
    phi = geoInvMlfn(y, g.en);
    double s = fabs(phi);
    if (s < kGeoHalfPi) {
      s = sin(phi);
      lam = x * sqrt(1. - g.projEs * s * s) / cos(phi);
    } else if ((s - 1e-10) < kGeoHalfPi) {
      lam = 0.;
    } else {
      lam = nan("");
      ok = false;
    }
  }
  lam = geoAdjlon(lam + g.lam0);
  return ok;
}

// Geographic radians -> map coordinates: fwd_prepare PROJ/src/fwd.cpp:84-113 (longitude
// relative to the central meridian), the projection's forward, fwd_finalize (scale by a,
// false easting / northing).
UB_DEV void geoProjForward(const GeoGrid &g, double lam, double phi, double &mx, double &my) {
  if (g.projection == USGSCSM_B200_PROJ_LONGLAT) {
    mx = lam * kGeoRadToDeg;
    my = phi * kGeoRadToDeg;
    return;
  }
  lam = geoAdjlon(lam);
  lam = geoAdjlon(lam - g.lam0);
  double x, y;
  if (g.projection == USGSCSM_B200_PROJ_EQC) {  // eqc.cpp:42-50
    x = g.rc * lam;
    y = phi - g.phi0;
  } else if (!g.ellipsoidal) {  // gn_sinu.cpp s_forward :54-83
    x = lam * cos(phi);
    y = phi;
  } else {  // gn_sinu.cpp e_forward :24-34
    double s = sin(phi), c = cos(phi);
    y = geoMlfn(phi, s, c, g.en);
    x = lam * c / sqrt(1. - g.projEs * s * s);
  }
  mx = g.projA * x + g.x0;
  my = g.projA * y + g.y0;
}

// Geodetic -> cartesian, PROJ/src/conversions/cart.cpp:107-143, from ready-made trigonometry.
UB_DEV void geoCartFromTrig(double N, double NoneEs, double cosphi, double sinphi, double coslam,
                            double sinlam, double h, double &x, double &y, double &z) {
  x = (N + h) * cosphi * coslam;
  y = (N + h) * cosphi * sinlam;
  z = (NoneEs + h) * sinphi;
}
UB_DEV void geoCartForward(const GeoGrid &g, double lam, double phi, double h, double &x,
                           double &y, double &z) {
  double sinphi, cosphi, sinlam, coslam;
  sincos(phi, &sinphi, &cosphi);
  sincos(lam, &sinlam, &coslam);
  const double N = g.es == 0 ? g.a : g.a / sqrt(1 - g.es * sinphi * sinphi);
  geoCartFromTrig(N, N * (1 - g.es), cosphi, sinphi, coslam, sinlam, h, x, y, z);
}

// Cartesian -> geodetic, PROJ/src/conversions/cart.cpp:146-199 (Bowring's closed form with the
// pole handling of the library).
UB_DEV void geoCartInverse(const GeoGrid &g, double X, double Y, double Z, double &lam, double &phi,
                           double &h) {
  const double p = hypot(X, Y);
  // second eccentricity squared: e2s = tan(asin(e))^2 (ell_set.cpp:585-586) == es / (1 - es)
  const double e2s = g.es / (1. - g.es);
  const double yTheta = Z * g.a;
  const double xTheta = p * g.b;
  const double norm = hypot(yTheta, xTheta);
  const double c = norm == 0 ? 1 : xTheta / norm;
  const double s = norm == 0 ? 0 : yTheta / norm;
  const double yPhi = Z + e2s * g.b * s * s * s;
  const double xPhi = p - g.es * g.a * c * c * c;
  const double normPhi = hypot(yPhi, xPhi);
  double cosphi = normPhi == 0 ? 1 : xPhi / normPhi;
  double sinphi = normPhi == 0 ? 0 : yPhi / normPhi;
  if (xPhi <= 0) {
    phi = Z >= 0 ? kGeoHalfPi : -kGeoHalfPi;
    cosphi = 0;
    sinphi = Z >= 0 ? 1 : -1;
  } else {
    phi = atan(yPhi / xPhi);
  }
  lam = atan2(Y, X);
  if (cosphi < 1e-6) {
    const double r = hypot(g.a * g.a * cosphi, g.b * g.b * sinphi) / hypot(g.a * cosphi, g.b * sinphi);
    h = fabs(Z) - r;
  } else {
    const double N = g.es == 0 ? g.a : g.a / sqrt(1 - g.es * sinphi * sinphi);
    h = p / cosphi - N;
  }
}

// pixelToMeter / meterToPixel, src/Utilities.cpp:2720-2745
UB_DEV void geoPixelToMeter(const GeoGrid &g, double line, double sample, double &mx, double &my) {
  mx = (sample * g.gt[1]) + g.gt[0];
  my = (line * g.gt[5]) + g.gt[3];
}
UB_DEV void geoMeterToPixel(const GeoGrid &g, double mx, double my, double &line, double &sample) {
  sample = (mx - g.gt[0]) / g.gt[1];
  line = (my - g.gt[3]) / g.gt[5];
}

// One raw DEM sample as a double; false when it is the nodata value.
UB_DEV bool demRaw(const void *samples, int type, long index, int hasNodata, double nodata,
                   double &raw) {
  if (type == USGSCSM_B200_DEM_F64) raw = __ldcs(static_cast<const double *>(samples) + index);
  else if (type == USGSCSM_B200_DEM_F32) raw = (double)__ldcs(static_cast<const float *>(samples) + index);
  else raw = (double)__ldcs(static_cast<const short *>(samples) + index);
  if (hasNodata && (raw == nodata || (isnan(nodata) && isnan(raw)))) return false;
  return true;
}

// Ground point of window pixel k = r * nCols + c.  False for a nodata sample.
UB_DEV bool demGroundPoint(const DemWindow &w, long k, double &x, double &y, double &z) {
  const long r = k / w.nCols;
  const long c = k - r * w.nCols;
  double raw;
  if (!demRaw(w.samples, w.sampleType, r * w.rowStride + c, w.hasNodata, w.nodata, raw)) {
    x = y = z = nan("");
    return false;
  }
  const double h = raw * w.scale + w.offset;
  if (w.rowTab) {
    const double2 *re = reinterpret_cast<const double2 *>(w.rowTab) + 2 * r;
    const double2 n = __ldg(re), t = __ldg(re + 1);
    const double2 ce = __ldg(reinterpret_cast<const double2 *>(w.colTab) + c);
    geoCartFromTrig(n.x, n.y, t.x, t.y, ce.x, ce.y, h, x, y, z);
  } else {
    double mx, my, lam, phi;
    geoPixelToMeter(w.grid, w.line0 + (double)r, w.samp0 + (double)c, mx, my);
    if (!geoProjInverse(w.grid, mx, my, lam, phi)) {
      x = y = z = nan("");
      return true;
    }
    geoCartForward(w.grid, lam, phi, h, x, y, z);
  }
  return true;
}

}  // namespace ub200
#endif
