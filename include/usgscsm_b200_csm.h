// usgscsm_b200_csm.h -- the engine as a REAL Community Sensor Model plugin.
//
// libusgscsm_b200_csm.so is compiled against the CSM 3.0.3 API headers (csm/*.h of the reference
// tree: the public interface, used as it is) and holds
//   * usgscsm_b200::UsgsAstroPlugin : ::csm::Plugin -- same plugin / manufacturer / model names as
//     the reference's (src/UsgsAstroPlugin.cpp:62-69), with the static self-registering instance
//     (src/UsgsAstroPlugin.cpp:72, csm/Plugin.cpp:36-117): a SET that dlopen()s the library finds it
//     with csm::Plugin::findPlugin("UsgsAstroPluginCSM");
//   * four ::csm::RasterGM (+ ::csm::SettableEllipsoid) subclasses, one per sensor model, whose
//     per-point virtuals forward to a batch of one through the C-ABI (usgscsm_b200.h) and which add
//     the batched calls the reference does not have (groundToImageBatch, imageToGroundBatch, ...).
// Install it INSTEAD of the reference's libusgscsm.so (both register the same plugin name).
// There is no CPU implementation behind it: without a CUDA device the constructors throw
// csm::Error(SENSOR_MODEL_NOT_CONSTRUCTIBLE).
#ifndef USGSCSM_B200_CSM_H_
#define USGSCSM_B200_CSM_H_

#include <string>
#include <vector>

#include "Plugin.h"
#include "RasterGM.h"
#include "CorrelationModel.h"
#include "SettableEllipsoid.h"

#include "usgscsm_b200.h"

namespace usgscsm_b200 {

// Common implementation of the four models; the subclasses supply the per-model metadata
// (names, units, sensor type ...) the reference hard-codes per class.
class SensorModel : public ::csm::RasterGM, virtual public ::csm::SettableEllipsoid {
 public:
  virtual ~SensorModel();

  // ---- new here: whole arrays per call (host pointers; SoA) ---------------------------------
  // status (nullable): USGSCSM_B200_ST_* bits per point instead of warnings / exceptions.
  void groundToImageBatch(size_t n, const double *x, const double *y, const double *z,
                          double desiredPrecision, double *line, double *samp,
                          double *achievedPrecision = NULL, uint32_t *status = NULL) const;
  void imageToGroundBatch(size_t n, const double *line, const double *samp, const double *height,
                          bool heightPerPoint, double desiredPrecision, double *x, double *y,
                          double *z, double *achievedPrecision = NULL, uint32_t *status = NULL) const;
  // arrays of the CSM value types (AoS)
  void groundToImageBatch(const std::vector< ::csm::EcefCoord> &groundPts, double desiredPrecision,
                          std::vector< ::csm::ImageCoord> &imagePts,
                          std::vector<uint32_t> *status = NULL) const;
  void imageToGroundBatch(const std::vector< ::csm::ImageCoord> &imagePts, double height,
                          double desiredPrecision, std::vector< ::csm::EcefCoord> &groundPts,
                          std::vector<uint32_t> *status = NULL) const;
  usgscsm_b200_model *handle() const { return m_h; }
  int device() const;

  // ---- csm::Model ------------------------------------------------------------------------------
  virtual ::csm::Version getVersion() const;
  virtual std::string getModelName() const;
  virtual std::string getImageIdentifier() const;
  virtual void setImageIdentifier(const std::string &imageId, ::csm::WarningList *warnings = NULL);
  virtual std::string getSensorIdentifier() const;
  virtual std::string getPlatformIdentifier() const;
  virtual std::string getCollectionIdentifier() const;
  virtual std::string getTrajectoryIdentifier() const;
  virtual std::string getReferenceDateAndTime() const;
  virtual std::string getModelState() const;
  virtual void replaceModelState(const std::string &argState);

  // ---- csm::GeometricModel ---------------------------------------------------------------------
  virtual ::csm::EcefCoord getReferencePoint() const;
  virtual void setReferencePoint(const ::csm::EcefCoord &groundPt);
  virtual int getNumParameters() const;
  virtual std::string getParameterUnits(int index) const;
  virtual bool hasShareableParameters() const;
  virtual bool isParameterShareable(int index) const;
  virtual ::csm::SharingCriteria getParameterSharingCriteria(int index) const;
  virtual double getParameterValue(int index) const;
  virtual void setParameterValue(int index, double value);
  virtual ::csm::param::Type getParameterType(int index) const;
  virtual void setParameterType(int index, ::csm::param::Type pType);
  virtual double getParameterCovariance(int index1, int index2) const;
  virtual void setParameterCovariance(int index1, int index2, double covariance);
  virtual int getNumGeometricCorrectionSwitches() const;
  virtual std::string getGeometricCorrectionName(int index) const;
  virtual void setGeometricCorrectionSwitch(int index, bool value, ::csm::param::Type pType);
  virtual bool getGeometricCorrectionSwitch(int index) const;
  virtual std::vector<double> getCrossCovarianceMatrix(
      const ::csm::GeometricModel &comparisonModel, ::csm::param::Set pSet = ::csm::param::VALID,
      const ::csm::GeometricModel::GeometricModelList &otherModels =
          ::csm::GeometricModel::GeometricModelList()) const;

  // ---- csm::RasterGM: one point per call = a batch of one on the GPU ------------------------------
  virtual ::csm::ImageCoord groundToImage(const ::csm::EcefCoord &groundPt,
                                          double desiredPrecision = 0.001,
                                          double *achievedPrecision = NULL,
                                          ::csm::WarningList *warnings = NULL) const;
  virtual ::csm::ImageCoordCovar groundToImage(const ::csm::EcefCoordCovar &groundPt,
                                               double desiredPrecision = 0.001,
                                               double *achievedPrecision = NULL,
                                               ::csm::WarningList *warnings = NULL) const;
  virtual ::csm::EcefCoord imageToGround(const ::csm::ImageCoord &imagePt, double height,
                                         double desiredPrecision = 0.001,
                                         double *achievedPrecision = NULL,
                                         ::csm::WarningList *warnings = NULL) const;
  virtual ::csm::EcefCoordCovar imageToGround(const ::csm::ImageCoordCovar &imagePt, double height,
                                              double heightVariance, double desiredPrecision = 0.001,
                                              double *achievedPrecision = NULL,
                                              ::csm::WarningList *warnings = NULL) const;
  virtual ::csm::EcefLocus imageToProximateImagingLocus(const ::csm::ImageCoord &imagePt,
                                                        const ::csm::EcefCoord &groundPt,
                                                        double desiredPrecision = 0.001,
                                                        double *achievedPrecision = NULL,
                                                        ::csm::WarningList *warnings = NULL) const;
  virtual ::csm::EcefLocus imageToRemoteImagingLocus(const ::csm::ImageCoord &imagePt,
                                                     double desiredPrecision = 0.001,
                                                     double *achievedPrecision = NULL,
                                                     ::csm::WarningList *warnings = NULL) const;
  virtual ::csm::ImageCoord getImageStart() const;
  virtual ::csm::ImageVector getImageSize() const;
  virtual std::pair< ::csm::ImageCoord, ::csm::ImageCoord> getValidImageRange() const;
  virtual std::pair<double, double> getValidHeightRange() const;
  virtual ::csm::EcefVector getIlluminationDirection(const ::csm::EcefCoord &groundPt) const;
  virtual double getImageTime(const ::csm::ImageCoord &imagePt) const;
  virtual ::csm::EcefCoord getSensorPosition(const ::csm::ImageCoord &imagePt) const;
  virtual ::csm::EcefCoord getSensorPosition(double time) const;
  virtual ::csm::EcefVector getSensorVelocity(const ::csm::ImageCoord &imagePt) const;
  virtual ::csm::EcefVector getSensorVelocity(double time) const;
  virtual ::csm::RasterGM::SensorPartials computeSensorPartials(
      int index, const ::csm::EcefCoord &groundPt, double desiredPrecision = 0.001,
      double *achievedPrecision = NULL, ::csm::WarningList *warnings = NULL) const;
  virtual ::csm::RasterGM::SensorPartials computeSensorPartials(
      int index, const ::csm::ImageCoord &imagePt, const ::csm::EcefCoord &groundPt,
      double desiredPrecision = 0.001, double *achievedPrecision = NULL,
      ::csm::WarningList *warnings = NULL) const;
  virtual std::vector< ::csm::RasterGM::SensorPartials> computeAllSensorPartials(
      const ::csm::EcefCoord &groundPt, ::csm::param::Set pSet = ::csm::param::VALID,
      double desiredPrecision = 0.001, double *achievedPrecision = NULL,
      ::csm::WarningList *warnings = NULL) const;
  virtual std::vector< ::csm::RasterGM::SensorPartials> computeAllSensorPartials(
      const ::csm::ImageCoord &imagePt, const ::csm::EcefCoord &groundPt,
      ::csm::param::Set pSet = ::csm::param::VALID, double desiredPrecision = 0.001,
      double *achievedPrecision = NULL, ::csm::WarningList *warnings = NULL) const;
  virtual std::vector<double> computeGroundPartials(const ::csm::EcefCoord &groundPt) const;
  virtual const ::csm::CorrelationModel &getCorrelationModel() const;
  virtual std::vector<double> getUnmodeledCrossCovariance(const ::csm::ImageCoord &pt1,
                                                          const ::csm::ImageCoord &pt2) const;

  // ---- csm::SettableEllipsoid -----------------------------------------------------------------------
  virtual ::csm::Ellipsoid getEllipsoid() const;
  virtual void setEllipsoid(const ::csm::Ellipsoid &ellipsoid);

 protected:
  SensorModel(const std::string &state, int device);
  void load(const std::string &state);
  std::string stateString(const char *key, const std::string &fallback) const;
  double stateNumber(const char *key, double fallback) const;

  usgscsm_b200_model *m_h;
  int m_device;
  struct Meta;   // parsed copy of the state: identifiers, height range, parameter types ...
  Meta *m_meta;
  ::csm::NoCorrelationModel m_noCorrelation;
};

class LineScannerModel : public SensorModel {  // UsgsAstroLsSensorModel
 public:
  LineScannerModel(const std::string &state, int device) : SensorModel(state, device) {}
  virtual std::string getPedigree() const;
  virtual std::string getSensorType() const;
  virtual std::string getSensorMode() const;
  virtual std::string getParameterName(int index) const;
};
class PushFrameModel : public SensorModel {  // UsgsAstroPushFrameSensorModel
 public:
  PushFrameModel(const std::string &state, int device) : SensorModel(state, device) {}
  virtual std::string getPedigree() const;
  virtual std::string getSensorType() const;
  virtual std::string getSensorMode() const;
  virtual std::string getParameterName(int index) const;
};
class FrameModel : public SensorModel {  // UsgsAstroFrameSensorModel
 public:
  FrameModel(const std::string &state, int device) : SensorModel(state, device) {}
  virtual std::string getPedigree() const;
  virtual std::string getSensorType() const;
  virtual std::string getSensorMode() const;
  virtual std::string getParameterName(int index) const;
  virtual std::string getParameterUnits(int index) const;
  virtual std::string getCollectionIdentifier() const;
  virtual std::string getTrajectoryIdentifier() const;
};
class SarModel : public SensorModel {  // UsgsAstroSarSensorModel
 public:
  SarModel(const std::string &state, int device) : SensorModel(state, device) {}
  virtual std::string getPedigree() const;
  virtual std::string getSensorType() const;
  virtual std::string getSensorMode() const;
  virtual std::string getParameterName(int index) const;
  virtual std::string getCollectionIdentifier() const;
  virtual std::string getTrajectoryIdentifier() const;
};

// include/usgscsm/UsgsAstroPlugin.h:11-85.  The CUDA device new models are placed on is
// USGSCSM_B200_DEVICE (default 0) or setDevice().
class UsgsAstroPlugin : public ::csm::Plugin {
 public:
  UsgsAstroPlugin();
  ~UsgsAstroPlugin();
  virtual std::string getPluginName() const;
  virtual std::string getManufacturer() const;
  virtual std::string getReleaseDate() const;
  virtual ::csm::Version getCsmVersion() const;
  virtual size_t getNumModels() const;
  virtual std::string getModelName(size_t modelIndex) const;
  virtual std::string getModelFamily(size_t modelIndex) const;
  virtual ::csm::Version getModelVersion(const std::string &modelName) const;
  virtual bool canModelBeConstructedFromState(const std::string &modelName,
                                              const std::string &modelState,
                                              ::csm::WarningList *warnings = NULL) const;
  virtual bool canModelBeConstructedFromISD(const ::csm::Isd &imageSupportData,
                                            const std::string &modelName,
                                            ::csm::WarningList *warnings = NULL) const;
  virtual ::csm::Model *constructModelFromState(const std::string &modelState,
                                                ::csm::WarningList *warnings = NULL) const;
  virtual ::csm::Model *constructModelFromISD(const ::csm::Isd &imageSupportData,
                                              const std::string &modelName,
                                              ::csm::WarningList *warnings = NULL) const;
  virtual std::string getModelNameFromModelState(const std::string &modelState,
                                                 ::csm::WarningList *warnings = NULL) const;
  virtual bool canISDBeConvertedToModelState(const ::csm::Isd &imageSupportData,
                                             const std::string &modelName,
                                             ::csm::WarningList *warnings = NULL) const;
  virtual std::string convertISDToModelState(const ::csm::Isd &imageSupportData,
                                             const std::string &modelName,
                                             ::csm::WarningList *warnings = NULL) const;
  std::string loadImageSupportData(const ::csm::Isd &imageSupportData, std::string *identifier) const;
  std::string getStateFromISD(::csm::Isd imageSupportData) const;
  void setDevice(int device) { m_device = device; }

 private:
  static const UsgsAstroPlugin m_registeredPlugin;
  int m_device;
};

}  // namespace usgscsm_b200
#endif
