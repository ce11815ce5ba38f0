// oracle/_ref driver: a C interface over the UNMODIFIED reference sensor models
// (TEST INFRASTRUCTURE -- not part of the product; only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load the library built from it).
//
// The reference sources are compiled where they lie under /root/reference by
// oracle/Makefile; this file only adds the loop a caller of the reference writes itself
// (/root/reference/bin/usgscsm_cam_test.cc:294-309): one csm::RasterGM virtual call per
// point, optionally under OpenMP over points (the calls are const and re-entrant,
// SURVEY.md §5).  Per point it records the result, achievedPrecision and a status word
// rebuilt from the csm::WarningList / thrown csm::Error, using the same bit layout as
// include/usgscsm_b200.h.
// Everything the model headers pull in is included first, so that the access hack below
// touches the four model class definitions only.
#include <any>
#include <sstream>
#include <nlohmann/json.hpp>
#include <RasterGM.h>
#include <CorrelationModel.h>
#include <SettableEllipsoid.h>
#include "Utilities.h"
#include "Distortion.h"
#include "EigenUtilities.h"
#include "VariantMap.h"
#include "ale/Orientations.h"
#include "ale/States.h"
#define private public  // oracle only: read m_projTransCoeffs & friends for the tests
#define protected public
#include "UsgsAstroFrameSensorModel.h"
#include "UsgsAstroLsSensorModel.h"
#include "UsgsAstroPushFrameSensorModel.h"
#include "UsgsAstroSarSensorModel.h"
#undef private
#undef protected
#include "UsgsAstroPlugin.h"
#include "Utilities.h"
#include "Distortion.h"
#include "EigenUtilities.h"

#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

enum : uint32_t {
  ST_NO_INTERSECTION = 1u << 0,
  ST_PRECISION_NOT_MET = 1u << 1,
  ST_ERROR_ALGORITHM = 1u << 5,
  ST_ERROR_UNKNOWN = 1u << 6,
  ST_NONFINITE = 1u << 7,
  ST_OUT_OF_IMAGE = 1u << 8,
  ST_ERROR_OTHER = 1u << 9,
};

struct RefModel {
  std::unique_ptr<csm::Model> model;
  csm::RasterGM *gm = nullptr;
  std::string state;
};

thread_local std::string g_lastError;

uint32_t bitsFromWarnings(const csm::WarningList &w) {
  uint32_t s = 0;
  for (const csm::Warning &x : w) {
    if (x.getWarning() == csm::Warning::NO_INTERSECTION) s |= ST_NO_INTERSECTION;
    if (x.getWarning() == csm::Warning::PRECISION_NOT_MET) s |= ST_PRECISION_NOT_MET;
  }
  return s;
}

uint32_t bitsFromError(const csm::Error &e) {
  if (e.getError() == csm::Error::ALGORITHM) return ST_ERROR_ALGORITHM;
  if (e.getError() == csm::Error::UNKNOWN_ERROR) return ST_ERROR_UNKNOWN;
  return ST_ERROR_OTHER;
}

int pickThreads(int threads) {
#ifdef _OPENMP
  return threads > 0 ? threads : omp_get_max_threads();
#else
  (void)threads;
  return 1;
#endif
}

}  // namespace

extern "C" {

const char *ref_last_error() { return g_lastError.c_str(); }

int ref_max_threads() { return pickThreads(0); }

// Build a model through the reference plugin from a model-state string
// ("<MODEL_NAME>\n{json}" or bare JSON; /root/reference/src/UsgsAstroPlugin.cpp:535-555).
void *ref_model_from_state(const char *state, size_t len) {
  try {
    UsgsAstroPlugin plugin;
    std::string s(state, len);
    csm::Model *m = plugin.constructModelFromState(s, nullptr);
    auto *rm = new RefModel;
    rm->model.reset(m);
    rm->gm = dynamic_cast<csm::RasterGM *>(m);
    if (!rm->gm) {
      delete rm;
      g_lastError = "model is not a csm::RasterGM";
      return nullptr;
    }
    return rm;
  } catch (const std::exception &e) {
    g_lastError = e.what();
    return nullptr;
  }
}

// Build a model through the reference plugin from an ISD, exactly as the reference's test
// fixtures do (/root/reference/tests/Fixtures.h:219-224): csm::Isd with the IMAGE file name,
// the plugin reads the sidecar "<basename>.json" (src/UsgsAstroPlugin.cpp:324-353, 488-519).
// model_name NULL / "" = the ISD's own "name_model" (UsgsAstroPlugin::getStateFromISD).
void *ref_model_from_isd(const char *image_path, const char *model_name) {
  try {
    UsgsAstroPlugin plugin;
    csm::Isd isd(image_path);
    std::string name = model_name ? model_name : "";
    if (name.empty()) {
      nlohmann::json j = nlohmann::json::parse(plugin.loadImageSupportData(isd));
      name = j.at("name_model").get<std::string>();
    }
    csm::Model *m = plugin.constructModelFromISD(isd, name, nullptr);
    auto *rm = new RefModel;
    rm->model.reset(m);
    rm->gm = dynamic_cast<csm::RasterGM *>(m);
    if (!rm->gm) {
      delete rm;
      g_lastError = "model is not a csm::RasterGM";
      return nullptr;
    }
    return rm;
  } catch (const std::exception &e) {
    g_lastError = e.what();
    return nullptr;
  }
}

// <Model>::constructStateFromIsd of an ISD text (what the plugin hands to populateModel, BEFORE
// the model derives gsd / reference point ...), serialised as JSON.  Returns the length; copies
// up to cap bytes.  0 = failure (ref_last_error).
size_t ref_state_from_isd(const char *isd_text, size_t len, char *out, size_t cap) {
  try {
    const std::string text(isd_text, len);
    nlohmann::json j = nlohmann::json::parse(text);
    const std::string name = j.at("name_model").get<std::string>();
    VariantMap vm;
    if (name == UsgsAstroFrameSensorModel::_SENSOR_MODEL_NAME) {
      UsgsAstroFrameSensorModel m;
      vm = m.constructStateFromIsd(text, nullptr);
    } else if (name == UsgsAstroLsSensorModel::_SENSOR_MODEL_NAME) {
      UsgsAstroLsSensorModel m;
      vm = m.constructStateFromIsd(text, nullptr);
    } else if (name == UsgsAstroPushFrameSensorModel::_SENSOR_MODEL_NAME) {
      UsgsAstroPushFrameSensorModel m;
      vm = m.constructStateFromIsd(text, nullptr);
    } else if (name == UsgsAstroSarSensorModel::_SENSOR_MODEL_NAME) {
      UsgsAstroSarSensorModel m;
      vm = m.constructStateFromIsd(text, nullptr);
    } else {
      g_lastError = "unsupported model name " + name;
      return 0;
    }
    const std::string s = jsonFromVariantMap(vm).dump();
    if (out && cap >= s.size()) std::memcpy(out, s.data(), s.size());
    return s.size();
  } catch (const std::exception &e) {
    g_lastError = e.what();
    return 0;
  }
}

void ref_model_free(void *h) { delete static_cast<RefModel *>(h); }

// Re-run the reference's own createProjectiveApproximation (LS:2745-2803), e.g. after
// ref_ls_set_projective replaced the seed.
int ref_ls_refit_projective(void *h) {
  RefModel *rm = static_cast<RefModel *>(h);
  auto *ls = dynamic_cast<UsgsAstroLsSensorModel *>(rm->model.get());
  if (!ls) return -1;
  try {
    ls->createProjectiveApproximation();
  } catch (...) {
    ls->m_useApproxInitTrans = false;
  }
  return 0;
}

// getModelState() of the live model; returns the length, copies up to cap bytes.
size_t ref_model_state(void *h, char *out, size_t cap) {
  RefModel *rm = static_cast<RefModel *>(h);
  rm->state = rm->model->getModelState();
  if (out && cap) {
    size_t n = rm->state.size() < cap ? rm->state.size() : cap;
    std::memcpy(out, rm->state.data(), n);
  }
  return rm->state.size();
}

// Host-derived quantities of a line-scan model that are inputs to the per-point math
// (SURVEY.md §3.1): out[0..3] = gsd, flyingHeight, halfSwath, halfTime; out[4] =
// useApproxInitTrans; out[5..18] = projTransCoeffs; out[19..21] = reference point.
int ref_ls_derived(void *h, double *out) {
  RefModel *rm = static_cast<RefModel *>(h);
  auto *ls = dynamic_cast<UsgsAstroLsSensorModel *>(rm->model.get());
  if (!ls) return -1;
  out[0] = ls->m_gsd;
  out[1] = ls->m_flyingHeight;
  out[2] = ls->m_halfSwath;
  out[3] = ls->m_halfTime;
  out[4] = ls->m_useApproxInitTrans ? 1.0 : 0.0;
  for (int i = 0; i < 14; i++)
    out[5 + i] = i < (int)ls->m_projTransCoeffs.size() ? ls->m_projTransCoeffs[i] : 0.0;
  out[19] = ls->m_referencePointXyz.x;
  out[20] = ls->m_referencePointXyz.y;
  out[21] = ls->m_referencePointXyz.z;
  return 0;
}

// Override the line-scan projective seed (so oracle and kernel can be fed identical
// host-fitted coefficients; the fit is an INPUT to the hot path).
int ref_ls_set_projective(void *h, int use, const double *coeffs14) {
  RefModel *rm = static_cast<RefModel *>(h);
  auto *ls = dynamic_cast<UsgsAstroLsSensorModel *>(rm->model.get());
  if (!ls) return -1;
  ls->m_useApproxInitTrans = use != 0;
  if (coeffs14) ls->m_projTransCoeffs.assign(coeffs14, coeffs14 + 14);
  return 0;
}

// Set an adjustable parameter through the CSM API (GeometricModel::setParameterValue).
int ref_set_parameter(void *h, int index, double value) {
  RefModel *rm = static_cast<RefModel *>(h);
  try {
    rm->gm->setParameterValue(index, value);
    return 0;
  } catch (const std::exception &e) {
    g_lastError = e.what();
    return -1;
  }
}

int ref_image_size(void *h, int *lines, int *samps) {
  RefModel *rm = static_cast<RefModel *>(h);
  csm::ImageVector v = rm->gm->getImageSize();
  *lines = (int)v.line;
  *samps = (int)v.samp;
  return 0;
}

// groundToImage over n points.  SoA inputs; outputs nullable except line/samp.
int ref_ground_to_image(void *h, long n, const double *x, const double *y, const double *z,
                        double prec, double *line, double *samp, double *achieved,
                        uint32_t *status, int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const auto range = gm->getValidImageRange();
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    double ach = std::nan("");
    csm::ImageCoord ip(std::nan(""), std::nan(""));
    try {
      csm::WarningList warn;
      ip = gm->groundToImage(csm::EcefCoord(x[i], y[i], z[i]), prec, &ach, &warn);
      st |= bitsFromWarnings(warn);
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
    }
    if (!(std::isfinite(ip.line) && std::isfinite(ip.samp))) st |= ST_NONFINITE;
    if (!(ip.line >= range.first.line && ip.line <= range.second.line &&
          ip.samp >= range.first.samp && ip.samp <= range.second.samp))
      st |= ST_OUT_OF_IMAGE;
    line[i] = ip.line;
    samp[i] = ip.samp;
    if (achieved) achieved[i] = ach;
    if (status) status[i] = st;
  }
  return 0;
}

// imageToGround over n points; `height` has n entries (height_stride 1) or 1 (stride 0).
int ref_image_to_ground(void *h, long n, const double *line, const double *samp,
                        const double *height, int height_stride, double prec, double *x,
                        double *y, double *z, double *achieved, uint32_t *status,
                        int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    double ach = std::nan("");
    csm::EcefCoord gp(std::nan(""), std::nan(""), std::nan(""));
    try {
      csm::WarningList warn;
      gp = gm->imageToGround(csm::ImageCoord(line[i], samp[i]),
                             height[height_stride ? i : 0], prec, &ach, &warn);
      st |= bitsFromWarnings(warn);
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
    }
    if (!(std::isfinite(gp.x) && std::isfinite(gp.y) && std::isfinite(gp.z))) st |= ST_NONFINITE;
    x[i] = gp.x;
    y[i] = gp.y;
    z[i] = gp.z;
    if (achieved) achieved[i] = ach;
    if (status) status[i] = st;
  }
  return 0;
}

// imageToRemoteImagingLocus over n pixels (csm::RasterGM virtual; LS:1254-1282 etc.).
int ref_remote_locus(void *h, long n, const double *line, const double *samp, double prec,
                     double *px, double *py, double *pz, double *dx, double *dy, double *dz,
                     double *achieved, uint32_t *status, int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    double ach = std::nan("");
    const double q = std::nan("");
    csm::EcefLocus loc(q, q, q, q, q, q);
    try {
      csm::WarningList warn;
      loc = gm->imageToRemoteImagingLocus(csm::ImageCoord(line[i], samp[i]), prec, &ach, &warn);
      st |= bitsFromWarnings(warn);
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
    }
    if (!(std::isfinite(loc.point.x) && std::isfinite(loc.point.y) && std::isfinite(loc.point.z) &&
          std::isfinite(loc.direction.x) && std::isfinite(loc.direction.y) &&
          std::isfinite(loc.direction.z)))
      st |= ST_NONFINITE;
    px[i] = loc.point.x;
    py[i] = loc.point.y;
    pz[i] = loc.point.z;
    dx[i] = loc.direction.x;
    dy[i] = loc.direction.y;
    dz[i] = loc.direction.z;
    if (achieved) achieved[i] = ach;
    if (status) status[i] = st;
  }
  return 0;
}

// imageToProximateImagingLocus over n (pixel, ground point) pairs.
int ref_proximate_locus(void *h, long n, const double *line, const double *samp, const double *x,
                        const double *y, const double *z, double prec, double *px, double *py,
                        double *pz, double *dx, double *dy, double *dz, double *achieved,
                        uint32_t *status, int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    double ach = std::nan("");
    const double q = std::nan("");
    csm::EcefLocus loc(q, q, q, q, q, q);
    try {
      csm::WarningList warn;
      loc = gm->imageToProximateImagingLocus(csm::ImageCoord(line[i], samp[i]),
                                             csm::EcefCoord(x[i], y[i], z[i]), prec, &ach, &warn);
      st |= bitsFromWarnings(warn);
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
      ach = q;
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
      ach = q;
    }
    if (!(std::isfinite(loc.point.x) && std::isfinite(loc.point.y) && std::isfinite(loc.point.z) &&
          std::isfinite(loc.direction.x) && std::isfinite(loc.direction.y) &&
          std::isfinite(loc.direction.z)))
      st |= ST_NONFINITE;
    px[i] = loc.point.x;
    py[i] = loc.point.y;
    pz[i] = loc.point.z;
    dx[i] = loc.direction.x;
    dy[i] = loc.direction.y;
    dz[i] = loc.direction.z;
    if (achieved) achieved[i] = ach;
    if (status) status[i] = st;
  }
  return 0;
}

// computeSensorPartials(index, ...) over n points; img_line / img_samp nullable (then the
// 5-argument overload computes the image point itself).
int ref_sensor_partials(void *h, int index, long n, const double *x, const double *y,
                        const double *z, const double *img_line, const double *img_samp,
                        double prec, double *dline, double *dsamp, double *achieved,
                        uint32_t *status, int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    double ach = std::nan("");
    csm::RasterGM::SensorPartials p(std::nan(""), std::nan(""));
    try {
      csm::WarningList warn;
      if (img_line && img_samp)
        p = gm->computeSensorPartials(index, csm::ImageCoord(img_line[i], img_samp[i]),
                                      csm::EcefCoord(x[i], y[i], z[i]), prec, &ach, &warn);
      else
        p = gm->computeSensorPartials(index, csm::EcefCoord(x[i], y[i], z[i]), prec, &ach, &warn);
      st |= bitsFromWarnings(warn);
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
      ach = std::nan("");
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
      ach = std::nan("");
    }
    if (!(std::isfinite(p.first) && std::isfinite(p.second))) st |= ST_NONFINITE;
    dline[i] = p.first;
    dsamp[i] = p.second;
    if (achieved) achieved[i] = ach;
    if (status) status[i] = st;
  }
  return 0;
}

// computeAllSensorPartials (pSet = VALID) over n points; partials = numParams * 2 doubles per
// point; img_line / img_samp nullable (then the ground-point overload).
int ref_all_sensor_partials(void *h, long n, const double *x, const double *y, const double *z,
                            const double *img_line, const double *img_samp, double prec,
                            double *partials, double *achieved, uint32_t *status, int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const int np = gm->getNumParameters();
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    double ach = std::nan("");
    std::vector<csm::RasterGM::SensorPartials> p;
    try {
      csm::WarningList warn;
      if (img_line && img_samp)
        p = gm->computeAllSensorPartials(csm::ImageCoord(img_line[i], img_samp[i]),
                                         csm::EcefCoord(x[i], y[i], z[i]), csm::param::VALID, prec,
                                         &ach, &warn);
      else
        p = gm->computeAllSensorPartials(csm::EcefCoord(x[i], y[i], z[i]), csm::param::VALID, prec,
                                         &ach, &warn);
      st |= bitsFromWarnings(warn);
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
      ach = std::nan("");
      p.clear();
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
      ach = std::nan("");
      p.clear();
    }
    for (int k = 0; k < np; k++) {
      const bool have = k < (int)p.size();
      const double a = have ? p[k].first : std::nan(""), b = have ? p[k].second : std::nan("");
      if (!(std::isfinite(a) && std::isfinite(b))) st |= ST_NONFINITE;
      partials[(i * np + k) * 2] = a;
      partials[(i * np + k) * 2 + 1] = b;
    }
    if (achieved) achieved[i] = ach;
    if (status) status[i] = st;
  }
  return np;
}

// getImageTime / getSensorPosition / getSensorVelocity per image point; out = 7 doubles per
// point (time, position, velocity).
int ref_image_sensor_state(void *h, long n, const double *line, const double *samp, double *out,
                           uint32_t *status, int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    double *o = out + i * 7;
    for (int k = 0; k < 7; k++) o[k] = std::nan("");
    const csm::ImageCoord ip(line[i], samp[i]);
    try {
      o[0] = gm->getImageTime(ip);
      const csm::EcefCoord p = gm->getSensorPosition(ip);
      o[1] = p.x;
      o[2] = p.y;
      o[3] = p.z;
      const csm::EcefVector v = gm->getSensorVelocity(ip);
      o[4] = v.x;
      o[5] = v.y;
      o[6] = v.z;
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
    }
    for (int k = 0; k < 7; k++)
      if (!std::isfinite(o[k])) st |= ST_NONFINITE;
    if (status) status[i] = st;
  }
  return 0;
}

// getIlluminationDirection per ground point; out = 3 doubles per point.
int ref_illumination_direction(void *h, long n, const double *x, const double *y, const double *z,
                               double *out, uint32_t *status, int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    double *o = out + i * 3;
    o[0] = o[1] = o[2] = std::nan("");
    try {
      const csm::EcefVector v = gm->getIlluminationDirection(csm::EcefCoord(x[i], y[i], z[i]));
      o[0] = v.x;
      o[1] = v.y;
      o[2] = v.z;
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
    }
    if (!(std::isfinite(o[0]) && std::isfinite(o[1]) && std::isfinite(o[2]))) st |= ST_NONFINITE;
    if (status) status[i] = st;
  }
  return 0;
}

// setParameterCovariance
int ref_set_parameter_covariance(void *h, int i, int j, double v) {
  RefModel *rm = static_cast<RefModel *>(h);
  rm->gm->setParameterCovariance(i, j, v);
  return 0;
}

// groundToImage(EcefCoordCovar): ground_cov 9 per point -> line, samp, image_cov 4 per point
int ref_ground_to_image_covar(void *h, long n, const double *x, const double *y, const double *z,
                              const double *ground_cov, double prec, double *line, double *samp,
                              double *image_cov, double *achieved, uint32_t *status, int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const std::pair<csm::ImageCoord, csm::ImageCoord> range = gm->getValidImageRange();
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    double ach = std::nan("");
    csm::ImageCoordCovar r(std::nan(""), std::nan(""));
    for (int k = 0; k < 4; k++) r.covariance[k] = std::nan("");
    try {
      csm::WarningList warn;
      csm::EcefCoordCovar g(x[i], y[i], z[i]);
      for (int k = 0; k < 9; k++) g.covariance[k] = ground_cov[i * 9 + k];
      r = gm->groundToImage(g, prec, &ach, &warn);
      st |= bitsFromWarnings(warn);
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
      ach = std::nan("");
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
      ach = std::nan("");
    }
    bool finite = true;
    for (int k = 0; k < 4; k++) {
      image_cov[i * 4 + k] = r.covariance[k];
      finite = finite && std::isfinite(r.covariance[k]);
    }
    if (!finite || !(std::isfinite(r.line) && std::isfinite(r.samp))) st |= ST_NONFINITE;
    if (!(r.line >= range.first.line && r.line <= range.second.line &&
          r.samp >= range.first.samp && r.samp <= range.second.samp))
      st |= ST_OUT_OF_IMAGE;  // derived, as in ref_ground_to_image
    line[i] = r.line;
    samp[i] = r.samp;
    if (achieved) achieved[i] = ach;
    if (status) status[i] = st;
  }
  return 0;
}

// imageToGround(ImageCoordCovar, height, heightVariance)
int ref_image_to_ground_covar(void *h, long n, const double *line, const double *samp,
                              const double *image_cov, const double *height, int height_stride,
                              const double *height_var, int var_stride, double prec, double *x,
                              double *y, double *z, double *ground_cov, double *achieved,
                              uint32_t *status, int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    double ach = std::nan("");
    csm::EcefCoordCovar r(std::nan(""), std::nan(""), std::nan(""));
    for (int k = 0; k < 9; k++) r.covariance[k] = std::nan("");
    try {
      csm::WarningList warn;
      csm::ImageCoordCovar p(line[i], samp[i]);
      for (int k = 0; k < 4; k++) p.covariance[k] = image_cov[i * 4 + k];
      r = gm->imageToGround(p, height[i * height_stride], height_var[i * var_stride], prec, &ach,
                            &warn);
      st |= bitsFromWarnings(warn);
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
      ach = std::nan("");
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
      ach = std::nan("");
    }
    bool finite = true;
    for (int k = 0; k < 9; k++) {
      ground_cov[i * 9 + k] = r.covariance[k];
      finite = finite && std::isfinite(r.covariance[k]);
    }
    if (!finite || !(std::isfinite(r.x) && std::isfinite(r.y) && std::isfinite(r.z))) st |= ST_NONFINITE;
    x[i] = r.x;
    y[i] = r.y;
    z[i] = r.z;
    if (achieved) achieved[i] = ach;
    if (status) status[i] = st;
  }
  return 0;
}

// computeGroundPartials over n points; partials = 6 doubles per point (LS:1299-1325 etc.).
int ref_ground_partials(void *h, long n, const double *x, const double *y, const double *z,
                        double *partials, uint32_t *status, int threads) {
  RefModel *rm = static_cast<RefModel *>(h);
  const csm::RasterGM *gm = rm->gm;
  const int nt = pickThreads(threads);
  (void)nt;
#pragma omp parallel for schedule(static) num_threads(nt)
  for (long i = 0; i < n; i++) {
    uint32_t st = 0;
    std::vector<double> p(6, std::nan(""));
    try {
      p = gm->computeGroundPartials(csm::EcefCoord(x[i], y[i], z[i]));
    } catch (const csm::Error &e) {
      st |= bitsFromError(e);
      p.assign(6, std::nan(""));
    } catch (const std::exception &) {
      st |= ST_ERROR_OTHER;
      p.assign(6, std::nan(""));
    }
    bool finite = true;
    for (int k = 0; k < 6; k++) {
      partials[6 * i + k] = p[k];
      finite = finite && std::isfinite(p[k]);
    }
    if (!finite) st |= ST_NONFINITE;
    if (status) status[i] = st;
  }
  return 0;
}

// nlohmann::json::to_msgpack of a model state (what bin/usgscsm_cam_test.cc
// --output-binary-model-state writes).  Returns the length; copies up to cap bytes.
size_t ref_state_to_msgpack(const char *state, size_t len, unsigned char *out, size_t cap) {
  try {
    nlohmann::json j = stateAsJson(std::string(state, len));
    std::vector<uint8_t> v = nlohmann::json::to_msgpack(j);
    if (out && cap >= v.size()) std::memcpy(out, v.data(), v.size());
    return v.size();
  } catch (const std::exception &e) {
    g_lastError = e.what();
    return 0;
  }
}

// csm::SettableEllipsoid::setEllipsoid (LS:2116-2119 etc.).
int ref_set_ellipsoid(void *h, double a, double b) {
  RefModel *rm = static_cast<RefModel *>(h);
  csm::Model *m = rm->model.get();
  const csm::Ellipsoid e(a, b);
  if (auto *ls = dynamic_cast<UsgsAstroLsSensorModel *>(m)) ls->setEllipsoid(e);
  else if (auto *pf = dynamic_cast<UsgsAstroPushFrameSensorModel *>(m)) pf->setEllipsoid(e);
  else if (auto *fr = dynamic_cast<UsgsAstroFrameSensorModel *>(m)) fr->setEllipsoid(e);
  else if (auto *sar = dynamic_cast<UsgsAstroSarSensorModel *>(m)) sar->setEllipsoid(e);
  else return -1;
  return 0;
}

// ---- L1 math entry points, for replaying the reference's known-answer tests ---------
// (/root/reference/tests/UtilitiesTests.cpp, DistortionTests.cpp, EigenUtilitiesTests.cpp)

int ref_lagrange(int numTime, const double *values, double startTime, double delTime,
                 double time, int vectorLength, int order, double *out) {
  try {
    lagrangeInterp(numTime, values, startTime, delTime, time, vectorLength, order, out);
    return 0;
  } catch (const std::exception &e) {
    g_lastError = e.what();
    return -1;
  }
}

int ref_remove_distortion(double dx, double dy, double *ux, double *uy, const double *coeffs,
                          int ncoeffs, double focal, int type, double tolerance) {
  try {
    std::vector<double> c(coeffs, coeffs + ncoeffs);
    removeDistortion(dx, dy, *ux, *uy, c, focal, (DistortionType)type, tolerance);
    return 0;
  } catch (const std::exception &e) {
    g_lastError = e.what();
    return -1;
  }
}

int ref_apply_distortion(double ux, double uy, double *dx, double *dy, const double *coeffs,
                         int ncoeffs, double focal, int type, double desiredPrecision,
                         double tolerance) {
  try {
    std::vector<double> c(coeffs, coeffs + ncoeffs);
    applyDistortion(ux, uy, *dx, *dy, c, focal, (DistortionType)type, desiredPrecision,
                    tolerance);
    return 0;
  } catch (const std::exception &e) {
    g_lastError = e.what();
    return -1;
  }
}

double ref_ellipsoid_elevation(double x, double y, double z, double a, double b, double prec,
                               double *achieved) {
  return computeEllipsoidElevation(x, y, z, a, b, prec, achieved);
}

void ref_rotation_from_quaternion(const double *q, double *m9) {
  double qq[4] = {q[0], q[1], q[2], q[3]};
  calculateRotationMatrixFromQuaternions(qq, m9);
}

void ref_rotation_from_euler(const double *e, double *m9) {
  double ee[3] = {e[0], e[1], e[2]};
  calculateRotationMatrixFromEuler(ee, m9);
}

int ref_projective_fit(int npts, const double *line, const double *samp, const double *x,
                       const double *y, const double *z, double *coeffs14) {
  try {
    std::vector<csm::ImageCoord> ip;
    std::vector<csm::EcefCoord> gp;
    for (int i = 0; i < npts; i++) {
      ip.push_back(csm::ImageCoord(line[i], samp[i]));
      gp.push_back(csm::EcefCoord(x[i], y[i], z[i]));
    }
    std::vector<double> u;
    usgscsm::computeBestFitProjectiveTransform(ip, gp, u);
    for (int i = 0; i < 14; i++) coeffs14[i] = u[i];
    return 0;
  } catch (const std::exception &e) {
    g_lastError = e.what();
    return -1;
  }
}

}  // extern "C"
