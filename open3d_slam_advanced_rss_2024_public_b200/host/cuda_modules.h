// host/cuda_modules.h — the registrable CUDA modules, by the names north_star gives them.  Each derives from the same
// abstract interface as its CPU twin, exposes the same static contract (description(), availableParameters(), ctor from
// Parameters — MatchersImpl.h:90-102 is the model) with the twin's parameter names / defaults / bounds plus `device`,
// and works standalone through the Eigen-shaped interface (upload -> kernel -> download) via the C ABI (include/b200reg.h).
// A chain made only of these modules never round-trips through host matrices: PM::ICP detects it and runs
// b200reg_icp_register (the "fast lane").
#pragma once
#include "../../include/b200reg.h"
#include "pm_compat.h"

namespace b200 {

// one C-ABI context per (thread, device); modules on the same thread + device share it (SURVEY 8b threading)
b200reg_ctx* contextFor(int device);
[[noreturn]] void throwStatus(b200reg_ctx* ctx, int32_t status);   // status code -> the libpointmatcher exception type

// Registers the CUDA modules into a PointMatcher<float> instance (the out-of-tree recipe of SURVEY 8b:
// const_cast<PM&>(PM::get()).MatcherRegistrar.reg("CudaKDTreeMatcher", new ...)).  Idempotent.
void registerCudaModules(PM& pm);

struct CudaVoxelGridDataPointsFilter : public PM::DataPointsFilter {   // twin: DataPointsFilters/VoxelGrid.h:66-75
  static const std::string description() { return "Construct Voxel grid of the point cloud (sm_100a). Down-sample by taking centroid or center of grid cells."; }
  static const PM::ParametersDoc availableParameters();
  const float vSizeX, vSizeY, vSizeZ;
  const bool useCentroid, averageExistingDescriptors;
  const int device;
  explicit CudaVoxelGridDataPointsFilter(const PM::Parameters& params = PM::Parameters());
  PM::DataPoints filter(const PM::DataPoints& input) override;
  void inPlaceFilter(PM::DataPoints& cloud) override;
};

struct CudaSurfaceNormalDataPointsFilter : public PM::DataPointsFilter {   // twin: DataPointsFilters/SurfaceNormal.h:66-82
  static const std::string description() { return "Surface normals and local statistics from the eigen-decomposition of the k nearest neighbours (sm_100a)."; }
  static const PM::ParametersDoc availableParameters();
  const unsigned knn;
  const float maxDist, epsilon;
  const bool keepNormals, keepDensities, keepEigenValues, keepEigenVectors, keepMatchedIds, keepMeanDist, keepSurfaceDescriptors, sortEigen, smoothNormals;
  const int device;
  explicit CudaSurfaceNormalDataPointsFilter(const PM::Parameters& params = PM::Parameters());
  PM::DataPoints filter(const PM::DataPoints& input) override;
  void inPlaceFilter(PM::DataPoints& cloud) override;
};

struct CudaKDTreeMatcher : public PM::Matcher {   // twin: MatchersImpl.h:88-112
  static const std::string description() { return "Matches a point from the reading to its closest neighbours in the reference: exact k-NN over a radix-sorted voxel grid (sm_100a)."; }
  static const PM::ParametersDoc availableParameters();
  const int knn;
  const float epsilon;
  const unsigned searchType;
  const float maxDist;
  const float cell;
  const int device;
  explicit CudaKDTreeMatcher(const PM::Parameters& params = PM::Parameters());
  ~CudaKDTreeMatcher() override;
  void init(const PM::DataPoints& filteredReference) override;
  PM::Matches findClosests(const PM::DataPoints& filteredReading) override;
 private:
  b200reg_index* index = nullptr;
};

struct CudaTrimmedDistOutlierFilter : public PM::OutlierFilter {   // twin: OutlierFiltersImpl.h:128-145
  static const std::string description() { return "Hard rejection threshold using quantile (radix select on the device)."; }
  static const PM::ParametersDoc availableParameters();
  const float ratio;
  const int device;
  explicit CudaTrimmedDistOutlierFilter(const PM::Parameters& params = PM::Parameters());
  PM::OutlierWeights compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference, const PM::Matches& input) override;
};

// MaxDist / MinDist / MedianDist twins (OutlierFiltersImpl.h:40-104): one device kernel behind all three (b200reg_dist_outlier)
struct CudaDistOutlierFilterBase : public PM::OutlierFilter {
  const int kind;      // B200REG_OUTLIER_*
  const float param;   // maxDist | minDist | factor
  const int device;
  CudaDistOutlierFilterBase(const std::string& name, const PM::ParametersDoc& doc, const PM::Parameters& params, int kind, const char* key);
  PM::OutlierWeights compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference, const PM::Matches& input) override;
};
struct CudaMaxDistOutlierFilter : public CudaDistOutlierFilterBase {
  static const std::string description() { return "This filter considers as outlier links whose norms are above a fix threshold."; }
  static const PM::ParametersDoc availableParameters();
  explicit CudaMaxDistOutlierFilter(const PM::Parameters& params = PM::Parameters());
};
struct CudaMinDistOutlierFilter : public CudaDistOutlierFilterBase {
  static const std::string description() { return "This filter considers as outlier links whose norms are below a threshold."; }
  static const PM::ParametersDoc availableParameters();
  explicit CudaMinDistOutlierFilter(const PM::Parameters& params = PM::Parameters());
};
struct CudaMedianDistOutlierFilter : public CudaDistOutlierFilterBase {
  static const std::string description() { return "This filter considers as outlier links whose norms are above the median link norms times a factor."; }
  static const PM::ParametersDoc availableParameters();
  explicit CudaMedianDistOutlierFilter(const PM::Parameters& params = PM::Parameters());
};

struct CudaVarTrimmedDistOutlierFilter : public PM::OutlierFilter {   // twin: OutlierFiltersImpl.h:147-173
  static const std::string description() { return "Hard rejection threshold using quantile and variable ratio. Based on \\cite{Phillips2007VarTrimmed}."; }
  static const PM::ParametersDoc availableParameters();
  const float minRatio, maxRatio, lambda;
  const int device;
  explicit CudaVarTrimmedDistOutlierFilter(const PM::Parameters& params = PM::Parameters());
  PM::OutlierWeights compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference, const PM::Matches& input) override;
};

struct CudaSurfaceNormalOutlierFilter : public PM::OutlierFilter {   // twin: OutlierFiltersImpl.h:176-192
  static const std::string description() { return "Hard rejection threshold using the angle between the surface normal vectors of the reading and the reference."; }
  static const PM::ParametersDoc availableParameters();
  const float maxAngle;
  const int device;
  explicit CudaSurfaceNormalOutlierFilter(const PM::Parameters& params = PM::Parameters());
  PM::OutlierWeights compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference, const PM::Matches& input) override;
};

struct CudaPointToPlaneErrorMinimizer : public PM::ErrorMinimizer {   // twin: ErrorMinimizers/PointToPlane.h:66-112
  static const std::string description() { return "Point-to-plane error (Chen & Medioni 1991); fused fp64 normal-equation reduction on the device."; }
  static const PM::ParametersDoc availableParameters();
  const bool force2D, force4DOF;
  const int device;
  explicit CudaPointToPlaneErrorMinimizer(const PM::Parameters& params = PM::Parameters());
  CudaPointToPlaneErrorMinimizer(const std::string& name, const PM::ParametersDoc& doc, const PM::Parameters& params);   // for the WithCov child
  PM::TransformationParameters compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference,
                                       const PM::OutlierWeights& outlierWeights, const PM::Matches& matches) override;
  // PointToPlane.cpp:410-422: sum of w ((reading - reference) . n)^2 over the ErrorElements pairs, reduced on the device
  float getResidualError(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference, const PM::OutlierWeights& outlierWeights,
                         const PM::Matches& matches) const override;
  // getOverlap (PointToPlane.cpp:426-521): DataPoints reaching this module carry no simpleSensorNoise / densities descriptors on the
  // device, the case for which the reference returns getWeightedPointUsedRatio() (:477-481) — the base class's implementation
 protected:
  // compute() with the pieces the covariance needs left on the device / returned: step transform, matched means
  PM::TransformationParameters computeOnDevice(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference,
                                               const PM::OutlierWeights& outlierWeights, const PM::Matches& matches, float sensorStdDev,
                                               PM::Matrix* covariance);
};

struct CudaPointToPointErrorMinimizer : public PM::ErrorMinimizer {   // twin: ErrorMinimizers/PointToPoint.h:45-62
  static const std::string description() { return "Point-to-point error. Based on SVD decomposition. Per \\cite{Besl1992Point2Point}."; }
  static const PM::ParametersDoc availableParameters();
  const int device;
  explicit CudaPointToPointErrorMinimizer(const PM::Parameters& params = PM::Parameters());
  PM::TransformationParameters compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference,
                                       const PM::OutlierWeights& outlierWeights, const PM::Matches& matches) override;
};

struct CudaPointToPlaneWithCovErrorMinimizer : public CudaPointToPlaneErrorMinimizer {   // twin: ErrorMinimizers/PointToPlaneWithCov.h:46-85
  static const std::string description() { return "Point-to-plane error (or point-to-line in 2D). Based on \\cite{Chen1991Point2Plane}. Covariance estimation based on \\cite{Censi2007ICPCovariance}."; }
  static const PM::ParametersDoc availableParameters();
  const float sensorStdDev;
  PM::Matrix covMatrix;
  explicit CudaPointToPlaneWithCovErrorMinimizer(const PM::Parameters& params = PM::Parameters());
  PM::TransformationParameters compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference,
                                       const PM::OutlierWeights& outlierWeights, const PM::Matches& matches) override;
  PM::Matrix getCovariance() const override { return covMatrix; }
};

// MaxDist / MinDist / BoundingBox / RandomSampling twins (DataPointsFilters/{MaxDist,MinDist,BoundingBox,RandomSampling}.h): one predicate
// kernel + an order-preserving compaction behind all four (b200reg_cloud_filter)
struct CudaCompactingFilterBase : public PM::DataPointsFilter {
  b200reg_filter_params fp;
  const int device;
  CudaCompactingFilterBase(const std::string& name, const PM::ParametersDoc& doc, const PM::Parameters& params);
  PM::DataPoints filter(const PM::DataPoints& input) override;
  void inPlaceFilter(PM::DataPoints& cloud) override;
};
struct CudaMaxDistDataPointsFilter : public CudaCompactingFilterBase {
  static const std::string description() { return "Subsampling. Filter points beyond a maximum distance measured on a specific axis. If dim is set to -1, points are filtered based on a maximum radius."; }
  static const PM::ParametersDoc availableParameters();
  explicit CudaMaxDistDataPointsFilter(const PM::Parameters& params = PM::Parameters());
};
struct CudaMinDistDataPointsFilter : public CudaCompactingFilterBase {
  static const std::string description() { return "Subsampling. Filter points before a minimum distance measured on a specific axis. If dim is set to -1, points are filtered based on a minimum radius."; }
  static const PM::ParametersDoc availableParameters();
  explicit CudaMinDistDataPointsFilter(const PM::Parameters& params = PM::Parameters());
};
struct CudaBoundingBoxDataPointsFilter : public CudaCompactingFilterBase {
  static const std::string description() { return "Subsampling. Remove points laying in a bounding box which is axis aligned."; }
  static const PM::ParametersDoc availableParameters();
  explicit CudaBoundingBoxDataPointsFilter(const PM::Parameters& params = PM::Parameters());
};
struct CudaRandomSamplingDataPointsFilter : public CudaCompactingFilterBase {
  static const std::string description() { return "Subsampling. This filter reduces the size of the point cloud by randomly dropping points. Based on \\cite{Masuda1996Random}"; }
  static const PM::ParametersDoc availableParameters();
  explicit CudaRandomSamplingDataPointsFilter(const PM::Parameters& params = PM::Parameters());
};

}  // namespace b200
