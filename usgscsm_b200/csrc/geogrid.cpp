// geogrid.cpp -- host side of the georeferenced-grid support (see geogrid.h): PROJ-string subset
// parser, ellipsoid constants the way PROJ derives them, and the per-row / per-column
// trigonometry tables of separable projections.
#include "geogrid.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <sstream>
#include <stdexcept>

namespace ub200 {

namespace {

constexpr double kDegToRad = 0.017453292519943296;  // PROJ/src/proj.h DEG_TO_RAD
constexpr double kPi = 3.14159265358979323846;
constexpr double kTwoPi = 6.28318530717958647693;

// Horner evaluation, PROJ/src/mlfn.cpp:12-18
double polyval(double x, const double *p, int N) {
  double y = N < 0 ? 0 : p[N];
  for (; N > 0;) y = y * x + p[--N];
  return y;
}

// pj_enfn, PROJ/src/mlfn.cpp:34-73: Fourier coefficients of the meridional distance in the third
// flattening n (6th order), both directions.
void meridionalCoefficients(double n, double en[kMlfnCoeffs]) {
  const int Lmax = 6;
  static const double coeffRad[] = {1, 1.0 / 4, 1.0 / 64, 1.0 / 256};
  static const double coeffMuPhi[] = {
      -3.0 / 2,    9.0 / 16,     -3.0 / 32,     15.0 / 16,    -15.0 / 32,    135.0 / 2048,
      -35.0 / 48,  105.0 / 256,  315.0 / 512,   -189.0 / 512, -693.0 / 1280, 1001.0 / 2048,
  };
  static const double coeffPhiMu[] = {
      3.0 / 2,     -27.0 / 32,   269.0 / 512,   21.0 / 16,        -55.0 / 32,     6759.0 / 4096,
      151.0 / 96,  -417.0 / 128, 1097.0 / 512,  -15543.0 / 2560,  8011.0 / 2560,  293393.0 / 61440,
  };
  double n2 = n * n, d = n;
  en[0] = polyval(n2, coeffRad, Lmax / 2) / (1 + n);
  for (int l = 0, o = 0; l < Lmax; ++l) {
    int m = (Lmax - l - 1) / 2;
    en[l + 1] = d * polyval(n2, coeffMuPhi + o, m);
    en[l + 1 + Lmax] = d * polyval(n2, coeffPhiMu + o, m);
    d *= n;
    o += m + 1;
  }
}

// Ellipsoid constants from (a, b): PROJ/src/ell_set.cpp:323-332 (the +b case) and
// pj_calc_ellipsoid_params :577-626.
void ellipsoidFromAxes(double a, double b, double &es, double &n) {
  if (b == a) {
    es = 0;
    n = 0;
    return;
  }
  const double f = (a - b) / a;
  es = 2 * f - f * f;
  const double e = std::sqrt(es);
  const double alpha = std::asin(e);
  n = std::pow(std::tan(alpha / 2), 2);
}

// PROJ/src/adjlon.cpp:7-23
double adjlon(double longitude) {
  if (std::fabs(longitude) < kPi + 1e-12) return longitude;
  longitude += kPi;
  longitude -= kTwoPi * std::floor(longitude / kTwoPi);
  longitude -= kPi;
  return longitude;
}

void finish(GeoGrid &g, double cartA, double cartB) {
  if (!(g.projA > 0)) throw std::invalid_argument("geogrid: the projection needs +R or +a");
  g.projRa = 1. / g.projA;
  g.a = cartA > 0 ? cartA : g.projA;
  g.b = cartB > 0 ? cartB : (cartA > 0 ? cartA : g.b);
  double n;
  ellipsoidFromAxes(g.a, g.b, g.es, n);
  g.oneEs = 1. - g.es;
  g.ra = 1. / g.a;
  if (g.projection == USGSCSM_B200_PROJ_EQC) {
    if (!(g.rc > 0)) throw std::invalid_argument("geogrid: eqc needs |lat_ts| < 90");  // eqc.cpp:84-88
  }
}

}  // namespace

GeoGrid geoGridFromC(const usgscsm_b200_geogrid &c) {
  GeoGrid g{};
  for (int i = 0; i < 6; i++) g.gt[i] = c.geotransform[i];
  if (c.geotransform[1] == 0.0 || c.geotransform[5] == 0.0)
    throw std::invalid_argument("geogrid: zero pixel size in the geotransform");
  g.projection = c.projection;
  if (c.projection != USGSCSM_B200_PROJ_LONGLAT && c.projection != USGSCSM_B200_PROJ_EQC &&
      c.projection != USGSCSM_B200_PROJ_SINU)
    throw std::invalid_argument("geogrid: unknown projection code");
  g.projA = c.proj_semi_major;
  double projB = c.proj_semi_minor > 0 ? c.proj_semi_minor : c.proj_semi_major;
  g.b = projB;
  double n = 0;
  ellipsoidFromAxes(g.projA, projB, g.projEs, n);
  g.x0 = c.x_0;
  g.y0 = c.y_0;
  g.lam0 = c.lon_0_deg * kDegToRad;
  g.phi0 = c.lat_0_deg * kDegToRad;
  g.rc = std::cos(c.lat_ts_deg * kDegToRad);
  if (c.projection == USGSCSM_B200_PROJ_EQC) g.projEs = 0;  // eqc.cpp:89: the sphere of radius a
  g.ellipsoidal = (c.projection == USGSCSM_B200_PROJ_SINU && g.projEs != 0.0) ? 1 : 0;
  for (double &v : g.en) v = 0;
  if (g.ellipsoidal) meridionalCoefficients(n, g.en);
  finish(g, c.semi_major, c.semi_minor);
  return g;
}

// "+key=value +flag ..." -> map; a bare "+key" maps to "".
static std::map<std::string, std::string> splitProj(const std::string &s) {
  std::map<std::string, std::string> kv;
  std::istringstream in(s);
  std::string tok;
  while (in >> tok) {
    if (tok[0] == '+') tok.erase(0, 1);
    if (tok.empty()) continue;
    const size_t eq = tok.find('=');
    if (eq == std::string::npos) kv[tok] = "";
    else kv[tok.substr(0, eq)] = tok.substr(eq + 1);
  }
  return kv;
}

usgscsm_b200_geogrid geoGridCFromProjString(const std::string &proj, const double gt[6],
                                            double cartA, double cartB) {
  usgscsm_b200_geogrid c;
  std::memset(&c, 0, sizeof(c));
  for (int i = 0; i < 6; i++) c.geotransform[i] = gt[i];
  std::map<std::string, std::string> kv = splitProj(proj);
  auto num = [&](const char *k, double def) {
    auto it = kv.find(k);
    if (it == kv.end()) return def;
    char *end = nullptr;
    double v = std::strtod(it->second.c_str(), &end);
    if (end == it->second.c_str()) throw std::invalid_argument(std::string("geogrid: bad value for +") + k);
    return v;
  };
  const std::string p = kv.count("proj") ? kv["proj"] : "";
  if (p == "longlat" || p == "latlong" || p == "lonlat" || p == "latlon")
    c.projection = USGSCSM_B200_PROJ_LONGLAT;
  else if (p == "eqc")
    c.projection = USGSCSM_B200_PROJ_EQC;
  else if (p == "sinu")
    c.projection = USGSCSM_B200_PROJ_SINU;
  else
    throw std::invalid_argument("geogrid: unsupported projection +proj=" + p +
                                " (longlat, eqc and sinu are built in)");
  static const char *known[] = {"proj", "lat_ts", "lat_0", "lon_0", "x_0", "y_0", "R", "a", "b",
                                "rf", "f", "units", "no_defs", "type", "over", "wktext"};
  for (auto &e : kv) {
    bool ok = false;
    for (const char *k : known) ok = ok || e.first == k;
    if (!ok) throw std::invalid_argument("geogrid: unsupported PROJ parameter +" + e.first);
  }
  if (kv.count("units") && kv["units"] != "m")
    throw std::invalid_argument("geogrid: only +units=m is supported");
  c.lat_ts_deg = num("lat_ts", 0.0);
  c.lat_0_deg = num("lat_0", 0.0);
  c.lon_0_deg = num("lon_0", 0.0);
  c.x_0 = num("x_0", 0.0);
  c.y_0 = num("y_0", 0.0);
  if (kv.count("R")) {
    c.proj_semi_major = c.proj_semi_minor = num("R", 0.0);
  } else if (kv.count("a")) {
    c.proj_semi_major = num("a", 0.0);
    if (kv.count("b")) c.proj_semi_minor = num("b", 0.0);
    else if (kv.count("rf")) c.proj_semi_minor = c.proj_semi_major * (1.0 - 1.0 / num("rf", 0.0));
    else if (kv.count("f")) c.proj_semi_minor = c.proj_semi_major * (1.0 - num("f", 0.0));
    else c.proj_semi_minor = c.proj_semi_major;
  } else if (cartA > 0) {
    // a geographic CRS without its own ellipsoid takes the cartesian step's
    c.proj_semi_major = cartA;
    c.proj_semi_minor = cartB > 0 ? cartB : cartA;
  } else {
    throw std::invalid_argument("geogrid: the PROJ string gives no ellipsoid (+R or +a [+b])");
  }
  c.semi_major = cartA;
  c.semi_minor = cartB;
  return c;
}

GeoGrid geoGridFromProjString(const std::string &proj, const double gt[6]) {
  return geoGridFromC(geoGridCFromProjString(proj, gt, 0.0, 0.0));
}

bool geoGridSeparable(const GeoGrid &g) {
  return g.projection == USGSCSM_B200_PROJ_LONGLAT || g.projection == USGSCSM_B200_PROJ_EQC;
}

void geoGridTables(const GeoGrid &g, double line0, double dline, int64_t nRows, double samp0,
                   double dsamp, int64_t nCols, std::vector<double> &rowTab,
                   std::vector<double> &colTab) {
  rowTab.resize((size_t)nRows * 4);
  colTab.resize((size_t)nCols * 2);
  for (int64_t r = 0; r < nRows; r++) {
    const double line = line0 + (double)r * dline;
    const double my = line * g.gt[5] + g.gt[3];  // pixelToMeter U:2722
    double phi;
    if (g.projection == USGSCSM_B200_PROJ_LONGLAT) {
      phi = my * kDegToRad;  // unitconvert deg -> rad
    } else {                 // eqc inverse: inv.cpp:41-57 then eqc.cpp:52-61
      const double y = (my - g.y0) * g.projRa;
      phi = y + g.phi0;
    }
    const double sinphi = std::sin(phi), cosphi = std::cos(phi);
    // normal_radius_of_curvature cart.cpp:107-113
    const double N = g.es == 0 ? g.a : g.a / std::sqrt(1 - g.es * sinphi * sinphi);
    double *e = &rowTab[(size_t)r * 4];
    e[0] = N;
    e[1] = N * (1 - g.es);
    e[2] = cosphi;
    e[3] = sinphi;
  }
  for (int64_t c = 0; c < nCols; c++) {
    const double samp = samp0 + (double)c * dsamp;
    const double mx = samp * g.gt[1] + g.gt[0];  // pixelToMeter U:2721
    double lam;
    if (g.projection == USGSCSM_B200_PROJ_LONGLAT) {
      lam = mx * kDegToRad;
    } else {
      const double x = (mx - g.x0) * g.projRa;
      lam = adjlon(x / g.rc + g.lam0);  // eqc.cpp:57, inv.cpp:113-121
    }
    colTab[(size_t)c * 2] = std::cos(lam);
    colTab[(size_t)c * 2 + 1] = std::sin(lam);
  }
}

}  // namespace ub200
