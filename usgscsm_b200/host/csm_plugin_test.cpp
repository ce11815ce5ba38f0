// csm_plugin_test.cpp -- a Sensor Exploitation Tool in miniature: it knows ONLY the CSM API
// (links libcsmapi, includes csm/*.h), dlopen()s a plugin library, finds the plugin by name with
// csm::Plugin::findPlugin and drives the model through the csm::RasterGM virtuals, one point per
// call, exactly as ISIS / ASP / SOCET GXP do.  Used by tests/test_csm_plugin_gpu.py to show that
// libusgscsm_b200_csm.so is a drop-in for the reference's libusgscsm.so.
//
//   csm_plugin_test <plugin.so> state <state file> <points.bin> <n> <out.bin>
//   csm_plugin_test <plugin.so> isd   <image file>  <points.bin> <n> <out.bin>
// points.bin: n x 3 doubles (line, samp, height), AoS.  out.bin: per point 10 doubles
// (x y z ach_i2g warn_i2g | line samp ach_g2i warn_g2i | error flag), then the model state text.
#include <dlfcn.h>

#include <cmath>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>
#include <vector>

#include "Error.h"
#include "Isd.h"
#include "Plugin.h"
#include "RasterGM.h"
#include "Warning.h"

static double warnBits(const csm::WarningList &w) {
  int b = 0;
  for (const csm::Warning &x : w) {
    if (x.getWarning() == csm::Warning::NO_INTERSECTION) b |= 1;
    if (x.getWarning() == csm::Warning::PRECISION_NOT_MET) b |= 2;
  }
  return b;
}

int main(int argc, char **argv) {
  if (argc != 7) {
    std::cerr << "usage: csm_plugin_test plugin.so state|isd file points.bin n out.bin\n";
    return 2;
  }
  if (!dlopen(argv[1], RTLD_NOW | RTLD_GLOBAL)) {
    std::cerr << "dlopen: " << dlerror() << "\n";
    return 1;
  }
  const csm::Plugin *plugin = csm::Plugin::findPlugin("UsgsAstroPluginCSM");
  if (!plugin) {
    std::cerr << "plugin UsgsAstroPluginCSM not registered\n";
    return 1;
  }
  std::cout << "plugin " << plugin->getPluginName() << " by " << plugin->getManufacturer() << ", "
            << plugin->getNumModels() << " models, CSM " << plugin->getCsmVersion().version() << "\n";
  std::unique_ptr<csm::Model> model;
  try {
    if (std::string(argv[2]) == "isd") {
      csm::Isd isd(argv[3]);
      std::ifstream f(std::string(argv[3]).substr(0, std::string(argv[3]).find_last_of('.')) + ".json");
      std::stringstream ss;
      ss << f.rdbuf();
      const std::string text = ss.str();
      const size_t p = text.find("\"USGS_ASTRO_");
      const std::string name = text.substr(p + 1, text.find('"', p + 1) - p - 1);
      if (!plugin->canModelBeConstructedFromISD(isd, name)) {
        std::cerr << "canModelBeConstructedFromISD is false\n";
        return 1;
      }
      model.reset(plugin->constructModelFromISD(isd, name));
    } else {
      std::ifstream f(argv[3], std::ios::binary);
      std::stringstream ss;
      ss << f.rdbuf();
      const std::string state = ss.str();
      const std::string name = plugin->getModelNameFromModelState(state);
      if (!plugin->canModelBeConstructedFromState(name, state)) {
        std::cerr << "canModelBeConstructedFromState is false\n";
        return 1;
      }
      model.reset(plugin->constructModelFromState(state));
    }
  } catch (const csm::Error &e) {
    std::cerr << "csm::Error " << e.getError() << ": " << e.getMessage() << " in " << e.getFunction() << "\n";
    return 1;
  }
  csm::RasterGM *gm = dynamic_cast<csm::RasterGM *>(model.get());
  if (!gm) {
    std::cerr << "not a csm::RasterGM\n";
    return 1;
  }
  // the line-scan / push-frame / SAR models project their reference point to date the image
  // (SAR:1732-1741); a state whose reference point does not project makes that call throw
  std::string date;
  try {
    date = gm->getReferenceDateAndTime();
  } catch (const csm::Error &e) {
    date = "<csm::Error: " + e.getMessage() + ">";
  }
  std::cout << "model " << gm->getModelName() << " family " << gm->getFamily() << " pedigree "
            << gm->getPedigree() << " image " << gm->getImageSize().line << " x "
            << gm->getImageSize().samp << " params " << gm->getNumParameters() << " ["
            << gm->getParameterName(0) << "] sensor " << gm->getSensorType() << "/" << gm->getSensorMode()
            << " id " << gm->getImageIdentifier() << " date " << date << "\n";
  const long n = std::atol(argv[5]);
  std::vector<double> pts((size_t)n * 3), out((size_t)n * 10);
  FILE *f = std::fopen(argv[4], "rb");
  if (!f || std::fread(pts.data(), sizeof(double), pts.size(), f) != pts.size()) return 1;
  std::fclose(f);
  const double nan = std::nan("");
  for (long i = 0; i < n; i++) {
    double *o = &out[(size_t)i * 10];
    for (int k = 0; k < 10; k++) o[k] = nan;
    o[9] = 0;
    try {
      csm::WarningList w1, w2;
      double a1 = nan, a2 = nan;
      const csm::EcefCoord g = gm->imageToGround(csm::ImageCoord(pts[3 * i], pts[3 * i + 1]), pts[3 * i + 2], 0.001, &a1, &w1);
      o[0] = g.x; o[1] = g.y; o[2] = g.z; o[3] = a1; o[4] = warnBits(w1);
      const csm::ImageCoord p = gm->groundToImage(g, 0.001, &a2, &w2);
      o[5] = p.line; o[6] = p.samp; o[7] = a2; o[8] = warnBits(w2);
    } catch (const csm::Error &e) {
      o[9] = 1 + (double)e.getError();
    }
  }
  f = std::fopen(argv[6], "wb");
  std::fwrite(out.data(), sizeof(double), out.size(), f);
  const std::string state = gm->getModelState();
  std::fwrite(state.data(), 1, state.size(), f);
  std::fclose(f);
  std::cout << "ok " << n << " points\n";
  return 0;
}
