// ALE shim (see Vectors.h): the ISD (JSON) accessors the reference calls from its
// constructStateFromIsd routines.  Signatures follow the call sites in
// /root/reference/src/UsgsAstro*SensorModel.cpp (e.g. UsgsAstroLsSensorModel.cpp:2827-3123).
#ifndef ORACLE_SHIM_ALE_UTIL_H
#define ORACLE_SHIM_ALE_UTIL_H
#include <string>
#include <vector>
#include <nlohmann/json.hpp>
#include "ale/Distortion.h"
#include "ale/Orientations.h"
#include "ale/States.h"
#include "ale/Vectors.h"
namespace ale {
using json = nlohmann::json;
std::string getSensorModelName(json isd);
std::string getImageId(json isd);
std::string getSensorName(json isd);
std::string getPlatformName(json isd);
std::string getIsisCameraVersion(json isd);
int getTotalLines(json isd);
int getTotalSamples(json isd);
double getStartingTime(json isd);
double getCenterTime(json isd);
std::vector<std::vector<double>> getLineScanRate(json isd);
int getSampleSumming(json isd);
int getLineSumming(json isd);
double getFocalLength(json isd);
double getFocalLengthUncertainty(json isd);
std::vector<double> getFocal2PixelLines(json isd);
std::vector<double> getFocal2PixelSamples(json isd);
double getDetectorCenterLine(json isd);
double getDetectorCenterSample(json isd);
double getDetectorStartingLine(json isd);
double getDetectorStartingSample(json isd);
double getMinHeight(json isd);
double getMaxHeight(json isd);
double getSemiMajorRadius(json isd);
double getSemiMinorRadius(json isd);
DistortionType getDistortionModel(json isd);
std::vector<double> getDistortionCoeffs(json isd);
std::vector<double> getSunPositions(json isd);
States getInstrumentPosition(json isd);
States getSunPosition(json isd);
Orientations getBodyRotation(json isd);
Orientations getInstrumentPointing(json isd);
std::vector<double> getGeoTransform(json isd);
std::string getProjection(json isd);
}  // namespace ale
#endif
