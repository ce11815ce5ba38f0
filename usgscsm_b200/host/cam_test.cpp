// usgscsm_b200_cam_test -- the reference's only user-facing tool for the hot path
// (bin/usgscsm_cam_test.cc, docs/source/tools/usgscsm_cam_test.rst) on top of the batched
// engine: load a camera model from a model-state file (.json state or GXP .sup), project every
// sample_rate-th pixel to the ground at a fixed height above the datum and back, and report
// the pixel errors.  The reference does this with one imageToGround and one groundToImage
// call per pixel (bin/usgscsm_cam_test.cc:294-309); here the two loops are two batch calls
// through include/usgscsm_b200_plugin.hpp (-> C-ABI -> sm_100a kernels).
//
// Same options, same report text.  Extra: --device N, --points-file (dump per-point results
// as hex doubles for parity checks), and a throughput line.  ISD input needs ALE and is not
// supported: the tool says so and exits 1.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "usgscsm_b200_plugin.hpp"

namespace ub = usgscsm_b200;

struct Options {
  std::string model, output_model_state, output_binary_state, points_file;
  int sample_rate, device;
  double subpixel_offset, height_above_datum, desired_precision;
  Options()
      : sample_rate(0), device(0), subpixel_offset(0.0), height_above_datum(0.0),
        desired_precision(0.0) {}
};

static void printUsage(std::string const &progName) {
  std::cout << "Usage: " << progName << " --model <model file> [other options]"
            << "\nSee the documentation for more information.\n";
}

// Every option starts with two dashes and has a string value (bin/usgscsm_cam_test.cc:47-119).
static bool parseOptions(int argc, char **argv, Options &opt) {
  std::vector<std::string> params;
  for (int it = 1; it < argc; it++) {
    std::string arg = argv[it];
    if (arg == "--help") {
      printUsage(argv[0]);
      return false;
    }
    params.push_back(arg);
  }
  if (params.size() % 2 != 0 || params.empty()) {
    std::cout << "Could not parse correctly the input arguments.\n";
    printUsage(argv[0]);
    return false;
  }
  std::map<std::string, std::string> parsed;
  for (size_t it = 0; it < params.size() / 2; it++) {
    std::string o = params[2 * it], v = params[2 * it + 1];
    if (o.size() <= 2 || o[0] != '-' || o[1] != '-' || v.empty() || v[0] == '-') {
      std::cout << "Could not parse correctly the input arguments.\n";
      printUsage(argv[0]);
      return false;
    }
    parsed[o.substr(2)] = v;
  }
  opt.model = parsed["model"];
  if (opt.model.empty()) {
    std::cout << "The input model file is empty.\n";
    printUsage(argv[0]);
    return false;
  }
  double rate = atof(parsed["sample-rate"].c_str());
  if (rate != std::round(rate)) {
    std::cout << "The value of --sample-rate must be an integer.\n";
    printUsage(argv[0]);
    return false;
  }
  if (parsed["desired-precision"].empty()) parsed["desired-precision"] = "0.001";
  opt.output_model_state = parsed["output-model-state"];
  opt.output_binary_state = parsed["output-binary-model-state"];
  opt.points_file = parsed["points-file"];
  opt.sample_rate = (int)rate;
  opt.device = atoi(parsed["device"].c_str());
  opt.subpixel_offset = atof(parsed["subpixel-offset"].c_str());
  opt.height_above_datum = atof(parsed["height-above-datum"].c_str());
  opt.desired_precision = atof(parsed["desired-precision"].c_str());
  return true;
}

// bin/usgscsm_cam_test.cc:122-141
static void printErrors(std::vector<double> &errors, double desired_precision,
                        double max_achieved_precision) {
  std::sort(errors.begin(), errors.end());
  if (errors.empty()) {
    std::cout << "Empty list of errors.\n";
    return;
  }
  std::cout << "Norm of pixel errors after projecting from camera to ground and back.\n";
  std::cout << "Min:    " << errors[0] << "\n";
  std::cout << "Median: " << errors[errors.size() / 2] << "\n";
  std::cout << "Max:    " << errors.back() << "\n";
  std::cout << "Count:  " << errors.size() << "\n";
  std::cout << std::endl;
  std::cout << "Desired precision: " << desired_precision << std::endl;
  std::cout << "Max achieved precision: " << max_achieved_precision << std::endl;
}

int main(int argc, char **argv) {
  Options opt;
  if (!parseOptions(argc, argv, opt)) return 1;

  std::ifstream ifs(opt.model.c_str(), std::ios::binary);
  if (!ifs) {
    std::cout << "Cannot open file: " << opt.model << std::endl;
    return 1;
  }
  std::stringstream buf;
  buf << ifs.rdbuf();
  std::string model_state = buf.str();

  ub::UsgsAstroPlugin plugin(opt.device);
  std::unique_ptr<ub::RasterGM> model;
  try {
    // isUsgsCsmIsd (src/Utilities.cpp:2977-3003): a JSON object with "body_rotation",
    // "instrument_position" and a "USGS_ASTRO_..." model name is an ISD
    const size_t first = model_state.find_first_not_of(" \t\n\r");
    const size_t nm = model_state.find("\"USGS_ASTRO_");
    if (first != std::string::npos && model_state[first] == '{' && nm != std::string::npos &&
        model_state.find("\"body_rotation\"") != std::string::npos &&
        model_state.find("\"instrument_position\"") != std::string::npos) {
      const std::string isd_model_name =
          model_state.substr(nm + 1, model_state.find("\"", nm + 1) - nm - 1);
      std::cout << "Detected JSON ISD with model: " << isd_model_name << "\n";
      model.reset(plugin.constructModelFromISD(opt.model, isd_model_name));
      std::cout << "Loaded a CSM model of type " << isd_model_name << " from ISD file " << opt.model
                << ".\n";
    } else {
    std::string name = plugin.getModelNameFromModelState(model_state);
    const unsigned char b0 = model_state.empty() ? 0 : (unsigned char)model_state[0];
    if ((b0 >= 0x80 && b0 <= 0x8F) || b0 == 0xDE || b0 == 0xDF)  // isMsgpack, U:2511-2530
      std::cout << "Detected msgpack binary model state: " << opt.model << "\n";
    std::cout << "Detected model state with model: " << name << "\n";
    model.reset(plugin.constructModelFromState(model_state));
    std::cout << "Loaded a CSM model of type " << name << " from model state file " << opt.model
              << ".\n";
    }
  } catch (const ub::csm::Error &e) {
    std::cerr << "Failed to load a CSM model from: " << opt.model << " (" << e.getMessage()
              << ").\n";
    return 1;
  }

  if (!opt.output_model_state.empty()) {
    std::cout << "Writing model state: " << opt.output_model_state << std::endl;
    std::ofstream ofs(opt.output_model_state.c_str());
    ofs << model->getModelState() << "\n";
  }

  if (!opt.output_binary_state.empty()) {  // bin/usgscsm_cam_test.cc --output-binary-model-state
    std::string state = model->getModelState();
    size_t need = 0;
    if (usgscsm_b200_state_to_msgpack(state.data(), state.size(), NULL, 0, &need) != 0) {
      std::cerr << usgscsm_b200_last_error() << "\n";
      return 1;
    }
    std::vector<unsigned char> bytes(need);
    usgscsm_b200_state_to_msgpack(state.data(), state.size(), bytes.data(), need, &need);
    std::cout << "Writing binary model state: " << opt.output_binary_state << " (" << need
              << " bytes)" << std::endl;
    std::ofstream ofs(opt.output_binary_state.c_str(), std::ios::binary);
    ofs.write(reinterpret_cast<const char *>(bytes.data()), (std::streamsize)need);
  }

  if (opt.sample_rate > 0) {
    ub::csm::ImageVector image_size = model->getImageSize();
    std::cout << "\n";
    std::cout << "Camera image rows and columns: " << image_size.line << ' ' << image_size.samp
              << "\n";
    std::cout << "Row and column sample rate: " << opt.sample_rate << "\n";
    std::cout << "Subpixel offset for each pixel: " << opt.subpixel_offset << "\n";
    std::cout << "Ground height (relative to datum): " << opt.height_above_datum << "\n";

    // the pixel list, in the reference's loop order (samples outer, lines inner)
    std::vector<ub::csm::ImageCoord> pix;
    for (int samp = 0; samp < image_size.samp; samp += opt.sample_rate)
      for (int line = 0; line < image_size.line; line += opt.sample_rate)
        pix.push_back(ub::csm::ImageCoord(line + opt.subpixel_offset, samp + opt.subpixel_offset));

    std::vector<ub::csm::EcefCoord> ground;
    std::vector<ub::csm::ImageCoord> back;
    std::vector<double> ach1, ach2, heights(1, opt.height_above_datum);
    std::vector<uint32_t> st1, st2;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    try {
      model->imageToGroundBatch(pix, heights, opt.desired_precision, ground, &ach1, &st1);
      model->groundToImageBatch(ground, opt.desired_precision, back, &ach2, &st2);
    } catch (const ub::csm::Error &e) {
      std::cerr << e.getFunction() << ": " << e.getMessage() << "\n";
      return 1;
    }
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    double max_achieved_precision = 0.0;
    std::vector<double> errors;
    errors.reserve(pix.size());
    for (size_t i = 0; i < pix.size(); i++) {
      // the scalar tool would have thrown at the first point with an error bit
      try {
        ub::RasterGM::raiseOrWarn(st1[i], NULL, "imageToGround");
        ub::RasterGM::raiseOrWarn(st2[i], NULL, "groundToImage");
      } catch (const ub::csm::Error &e) {
        std::cerr << e.getFunction() << ": " << e.getMessage() << " (pixel " << pix[i].line << ' '
                  << pix[i].samp << ")\n";
        return 1;
      }
      max_achieved_precision = std::max(max_achieved_precision, ach1[i]);
      max_achieved_precision = std::max(max_achieved_precision, ach2[i]);
      double dl = pix[i].line - back[i].line, ds = pix[i].samp - back[i].samp;
      errors.push_back(std::sqrt(dl * dl + ds * ds));
    }
    if (!opt.points_file.empty()) {
      FILE *f = fopen(opt.points_file.c_str(), "w");
      if (!f) {
        std::cerr << "Cannot write " << opt.points_file << "\n";
        return 1;
      }
      for (size_t i = 0; i < pix.size(); i++)
        fprintf(f, "%a %a %a %a %a %a %a %a %a %u %u\n", pix[i].line, pix[i].samp, ground[i].x,
                ground[i].y, ground[i].z, ach1[i], back[i].line, back[i].samp, ach2[i],
                st1[i] & USGSCSM_B200_ST_REFERENCE_MASK, st2[i] & USGSCSM_B200_ST_REFERENCE_MASK);
      fclose(f);
    }
    printErrors(errors, opt.desired_precision, max_achieved_precision);
    std::cout << "Batched on CUDA device " << model->device() << ": " << pix.size()
              << " pixels to ground and back in " << sec << " s ("
              << 2.0 * pix.size() / sec / 1e6 << " Mpoints/s, host arrays, copies included)\n";
  }
  return 0;
}
