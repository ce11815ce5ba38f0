/* oracle/geo_restate.c -- TEST INFRASTRUCTURE (parity checker), not part of the product.
 *
 * CPU restatement, in plain C, of the arithmetic UsgsAstroProjectedSensorModel delegates to PROJ
 * (src/UsgsAstroProjectedSensorModel.cpp:289-365: proj_trans between "+proj=<map projection>"
 * and "+proj=geocent +a +b", :128-147) plus pixelToMeter / meterToPixel
 * (src/Utilities.cpp:2720-2745), for the three projections the engine evaluates itself.
 * PROJ is vendored with the reference (PROJ/, release 9.x) but its library needs cmake, sqlite3
 * and generated sources to build, so it is "unbuildable" under this task's rules; each function
 * follows the PROJ source file:line it names.  Pinned by tests/test_oracle_geo.py on
 *   - PROJ's own regression vectors: PROJ/test/gie/more_builtins.gie:828-882 (cart, GRS80),
 *     PROJ/test/gie/builtins.gie:1359-1379 (eqc, a = 6400000);
 *   - the reference's golden values: tests/ProjectedCameraTests.cpp:32-58 (imageToGround of the
 *     constVelocityProj fixture, 1e-9), tests/UtilitiesTests.cpp:656-668 (pixelToMeter /
 *     meterToPixel).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this.
 */
#include <math.h>
#include <stdint.h>

#define PROJ_LONGLAT 0
#define PROJ_EQC 1
#define PROJ_SINU 2

#define DEG_TO_RAD .017453292519943296 /* PROJ/src/proj.h */
#define RAD_TO_DEG 57.295779513082321
#define M_HALFPI 1.57079632679489661923
#define M_TWOPI 6.28318530717958647693
#define LMAX 6

typedef struct {
  double gt[6];
  int projection;
  int ellipsoidal;
  double proj_a, proj_es; /* P->a, P->es of the projected CRS */
  double x0, y0, lam0, phi0, rc;
  double en[2 * LMAX + 1];
  double a, b, es; /* ellipsoid of the cart step */
} geo_t;

/* PROJ/src/mlfn.cpp:12-18 */
static double polyval(double x, const double p[], int N) {
  double y = N < 0 ? 0 : p[N];
  for (; N > 0;) y = y * x + p[--N];
  return y;
}
/* PROJ/src/mlfn.cpp:21-32 */
static double clenshaw(double szeta, double czeta, const double c[], int K) {
  double u0 = 0, u1 = 0, X = 2 * (czeta - szeta) * (czeta + szeta);
  for (; K > 0;) {
    double t = X * u0 - u1 + c[--K];
    u1 = u0;
    u0 = t;
  }
  return 2 * szeta * czeta * u0;
}
/* pj_enfn, PROJ/src/mlfn.cpp:34-73 */
static void enfn(double n, double *en) {
  static const double coeff_rad[] = {1, 1.0 / 4, 1.0 / 64, 1.0 / 256};
  static const double coeff_mu_phi[] = {
      -3.0 / 2,   9.0 / 16,    -3.0 / 32,   15.0 / 16,    -15.0 / 32,    135.0 / 2048,
      -35.0 / 48, 105.0 / 256, 315.0 / 512, -189.0 / 512, -693.0 / 1280, 1001.0 / 2048,
  };
  static const double coeff_phi_mu[] = {
      3.0 / 2,    -27.0 / 32,   269.0 / 512,  21.0 / 16,       -55.0 / 32,    6759.0 / 4096,
      151.0 / 96, -417.0 / 128, 1097.0 / 512, -15543.0 / 2560, 8011.0 / 2560, 293393.0 / 61440,
  };
  double n2 = n * n, d = n;
  en[0] = polyval(n2, coeff_rad, LMAX / 2) / (1 + n);
  for (int l = 0, o = 0; l < LMAX; ++l) {
    int m = (LMAX - l - 1) / 2;
    en[l + 1] = d * polyval(n2, coeff_mu_phi + o, m);
    en[l + 1 + LMAX] = d * polyval(n2, coeff_phi_mu + o, m);
    d *= n;
    o += m + 1;
  }
}
/* PROJ/src/mlfn.cpp:75-83 */
static double mlfn(double phi, double sphi, double cphi, const double *en) {
  return en[0] * (phi + clenshaw(sphi, cphi, en + 1, LMAX));
}
static double inv_mlfn(double mu, const double *en) {
  mu /= en[0];
  return mu + clenshaw(sin(mu), cos(mu), en + 1 + LMAX, LMAX);
}

/* PROJ/src/adjlon.cpp:7-23 */
static double adjlon(double longitude) {
  if (fabs(longitude) < M_PI + 1e-12) return longitude;
  longitude += M_PI;
  longitude -= M_TWOPI * floor(longitude / M_TWOPI);
  longitude -= M_PI;
  return longitude;
}

/* es and third flattening from the two axes: PROJ/src/ell_set.cpp:323-332 (+b), :577-610 */
static void ellipsoid(double a, double b, double *es, double *n) {
  if (a == b) {
    *es = 0;
    *n = 0;
    return;
  }
  double f = (a - b) / a;
  *es = 2 * f - f * f;
  double alpha = asin(sqrt(*es));
  *n = pow(tan(alpha / 2), 2);
}

/* params: gt[6], projection, lat_ts_deg, lat_0_deg, lon_0_deg, x_0, y_0, proj_a, proj_b,
 * cart_a, cart_b (0 = the projection's) */
void geo_setup(const double *gt, int projection, double lat_ts_deg, double lat_0_deg,
               double lon_0_deg, double x_0, double y_0, double proj_a, double proj_b,
               double cart_a, double cart_b, geo_t *g) {
  double n;
  for (int i = 0; i < 6; i++) g->gt[i] = gt[i];
  g->projection = projection;
  if (!(proj_b > 0)) proj_b = proj_a;
  g->proj_a = proj_a;
  ellipsoid(proj_a, proj_b, &g->proj_es, &n);
  if (projection == PROJ_EQC) g->proj_es = 0; /* PROJ/src/projections/eqc.cpp:89 */
  g->ellipsoidal = projection == PROJ_SINU && g->proj_es != 0;
  for (int i = 0; i < 2 * LMAX + 1; i++) g->en[i] = 0;
  if (g->ellipsoidal) enfn(n, g->en);
  g->x0 = x_0;
  g->y0 = y_0;
  g->lam0 = lon_0_deg * DEG_TO_RAD;
  g->phi0 = lat_0_deg * DEG_TO_RAD;
  g->rc = cos(lat_ts_deg * DEG_TO_RAD);
  g->a = cart_a > 0 ? cart_a : proj_a;
  g->b = cart_b > 0 ? cart_b : (cart_a > 0 ? cart_a : proj_b);
  ellipsoid(g->a, g->b, &g->es, &n);
}

/* map coordinates -> geographic radians: inv_prepare PROJ/src/inv.cpp:77-97, projection inverse
 * (eqc.cpp:52-61; gn_sinu.cpp:36-52 ellipsoid, :87-97 sphere), inv_finalize inv.cpp:113-121 */
static int proj_inverse(const geo_t *g, double mx, double my, double *lam, double *phi) {
  if (g->projection == PROJ_LONGLAT) {
    *lam = mx * DEG_TO_RAD;
    *phi = my * DEG_TO_RAD;
    return 1;
  }
  double ra = 1. / g->proj_a;
  double x = (mx - g->x0) * ra, y = (my - g->y0) * ra;
  int ok = 1;
  if (g->projection == PROJ_EQC) {
    *lam = x / g->rc;
    *phi = y + g->phi0;
  } else if (!g->ellipsoidal) {
    *phi = y;
    *lam = x / cos(y);
  } else {
    *phi = inv_mlfn(y, g->en);
    double s = fabs(*phi);
    if (s < M_HALFPI) {
      s = sin(*phi);
      *lam = x * sqrt(1. - g->proj_es * s * s) / cos(*phi);
    } else if ((s - 1e-10) < M_HALFPI) {
      *lam = 0.;
    } else {
      *lam = NAN;
      ok = 0;
    }
  }
  *lam = adjlon(*lam + g->lam0);
  return ok;
}

/* geographic radians -> map coordinates: fwd_prepare PROJ/src/fwd.cpp:84-113, projection
 * forward (eqc.cpp:42-50; gn_sinu.cpp:24-34, :54-83), fwd_finalize */
static void proj_forward(const geo_t *g, double lam, double phi, double *mx, double *my) {
  if (g->projection == PROJ_LONGLAT) {
    *mx = lam * RAD_TO_DEG;
    *my = phi * RAD_TO_DEG;
    return;
  }
  lam = adjlon(lam);
  lam = adjlon(lam - g->lam0);
  double x, y;
  if (g->projection == PROJ_EQC) {
    x = g->rc * lam;
    y = phi - g->phi0;
  } else if (!g->ellipsoidal) {
    x = lam * cos(phi);
    y = phi;
  } else {
    double s = sin(phi), c = cos(phi);
    y = mlfn(phi, s, c, g->en);
    x = lam * c / sqrt(1. - g->proj_es * s * s);
  }
  *mx = g->proj_a * x + g->x0;
  *my = g->proj_a * y + g->y0;
}

/* PROJ/src/conversions/cart.cpp:107-143 */
static void cart_forward(const geo_t *g, double lam, double phi, double h, double *xyz) {
  const double cosphi = cos(phi), sinphi = sin(phi);
  const double N = g->es == 0 ? g->a : g->a / sqrt(1 - g->es * sinphi * sinphi);
  xyz[0] = (N + h) * cosphi * cos(lam);
  xyz[1] = (N + h) * cosphi * sin(lam);
  xyz[2] = (N * (1 - g->es) + h) * sinphi;
}

/* PROJ/src/conversions/cart.cpp:146-199 */
static void cart_inverse(const geo_t *g, const double *xyz, double *lam, double *phi, double *h) {
  const double X = xyz[0], Y = xyz[1], Z = xyz[2];
  const double p = hypot(X, Y);
  const double e2 = tan(asin(sqrt(g->es))); /* ell_set.cpp:585-586 */
  const double e2s = e2 * e2;
  const double y_theta = Z * g->a, x_theta = p * g->b;
  const double norm = hypot(y_theta, x_theta);
  const double c = norm == 0 ? 1 : x_theta / norm;
  const double s = norm == 0 ? 0 : y_theta / norm;
  const double y_phi = Z + e2s * g->b * s * s * s;
  const double x_phi = p - g->es * g->a * c * c * c;
  const double norm_phi = hypot(y_phi, x_phi);
  double cosphi = norm_phi == 0 ? 1 : x_phi / norm_phi;
  double sinphi = norm_phi == 0 ? 0 : y_phi / norm_phi;
  if (x_phi <= 0) {
    *phi = Z >= 0 ? M_HALFPI : -M_HALFPI;
    cosphi = 0;
    sinphi = Z >= 0 ? 1 : -1;
  } else {
    *phi = atan(y_phi / x_phi);
  }
  *lam = atan2(Y, X);
  if (cosphi < 1e-6) {
    const double r = hypot(g->a * g->a * cosphi, g->b * g->b * sinphi) / hypot(g->a * cosphi, g->b * sinphi);
    *h = fabs(Z) - r;
  } else {
    const double N = g->es == 0 ? g->a : g->a / sqrt(1 - g->es * sinphi * sinphi);
    *h = p / cosphi - N;
  }
}

/* UsgsAstroProjectedSensorModel::imageToGround (src/UsgsAstroProjectedSensorModel.cpp:329-365)
 * with pixelToMeter (src/Utilities.cpp:2720-2725); height[i] (nullable: 0). */
void geo_image_to_ground(const geo_t *g, int64_t n, const double *line, const double *samp,
                         const double *height, double *x, double *y, double *z) {
  for (int64_t i = 0; i < n; i++) {
    double mx = (samp[i] * g->gt[1]) + g->gt[0];
    double my = (line[i] * g->gt[5]) + g->gt[3];
    double lam, phi, xyz[3] = {NAN, NAN, NAN};
    if (proj_inverse(g, mx, my, &lam, &phi)) cart_forward(g, lam, phi, height ? height[i] : 0.0, xyz);
    x[i] = xyz[0];
    y[i] = xyz[1];
    z[i] = xyz[2];
  }
}

/* UsgsAstroProjectedSensorModel::groundToImage (:289-321) with meterToPixel (U:2740-2745) */
void geo_ground_to_image(const geo_t *g, int64_t n, const double *x, const double *y,
                         const double *z, double *line, double *samp, double *height) {
  for (int64_t i = 0; i < n; i++) {
    double xyz[3] = {x[i], y[i], z[i]}, lam, phi, h, mx, my;
    cart_inverse(g, xyz, &lam, &phi, &h);
    proj_forward(g, lam, phi, &mx, &my);
    samp[i] = (mx - g->gt[0]) / g->gt[1];
    line[i] = (my - g->gt[3]) / g->gt[5];
    if (height) height[i] = h;
  }
}

/* single steps in PROJ's own units, for replaying its .gie vectors: degrees in / out */
void geo_cart_forward_deg(const geo_t *g, double lon_deg, double lat_deg, double h, double *xyz) {
  cart_forward(g, lon_deg * DEG_TO_RAD, lat_deg * DEG_TO_RAD, h, xyz);
}
void geo_cart_inverse_deg(const geo_t *g, const double *xyz, double *out3) {
  double lam, phi, h;
  cart_inverse(g, xyz, &lam, &phi, &h);
  out3[0] = lam * RAD_TO_DEG;
  out3[1] = phi * RAD_TO_DEG;
  out3[2] = h;
}
void geo_proj_forward_deg(const geo_t *g, double lon_deg, double lat_deg, double *xy) {
  proj_forward(g, lon_deg * DEG_TO_RAD, lat_deg * DEG_TO_RAD, &xy[0], &xy[1]);
}
void geo_proj_inverse_deg(const geo_t *g, double mx, double my, double *lonlat) {
  double lam, phi;
  proj_inverse(g, mx, my, &lam, &phi);
  lonlat[0] = lam * RAD_TO_DEG;
  lonlat[1] = phi * RAD_TO_DEG;
}
int geo_sizeof(void) { return (int)sizeof(geo_t); }
