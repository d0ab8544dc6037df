// host/cuda_modules.cpp — the registrable CUDA modules: thin C++ classes over the C ABI (include/b200reg.h).
// Host matrices cross the boundary as plain pointers in Eigen's column-major layout; status codes come back and are
// re-thrown as the libpointmatcher exception types callers already catch (Mapper.cpp:420 catches std::runtime_error).
#include "cuda_modules.h"

#include <cstring>
#include <map>
#include <mutex>
#include <thread>

namespace b200 {

typedef PM::Parametrizable P;

// ---------------------------------------------------------------------------------------------------------------
// contexts + error translation
// ---------------------------------------------------------------------------------------------------------------
b200reg_ctx* contextFor(int device) {
  struct Holder {
    std::map<int, b200reg_ctx*> ctxs;
    ~Holder() { for (auto& kv : ctxs) b200reg_ctx_destroy(kv.second); }
  };
  thread_local Holder holder;   // one ctx per (thread, device): ctxs are not thread-safe, distinct ctxs are independent
  auto it = holder.ctxs.find(device);
  if (it != holder.ctxs.end()) return it->second;
  b200reg_ctx* ctx = nullptr;
  const int32_t st = b200reg_ctx_create(device, &ctx);
  if (st != B200REG_OK) throw std::runtime_error("b200reg: no usable CUDA device " + std::to_string(device) + " (there is no CPU fallback)");
  holder.ctxs[device] = ctx;
  return ctx;
}

void throwStatus(b200reg_ctx* ctx, int32_t status) {
  const std::string msg = b200reg_last_error(ctx);
  switch (status) {
    case B200REG_E_INVALID_PARAM: throw PM::InvalidParameter(msg);
    case B200REG_E_INVALID_FIELD: throw PM::DataPoints::InvalidField(msg);
    case B200REG_E_NOT_CONVERGED: throw PM::ConvergenceError(msg);
    case B200REG_E_TRANSFORM: throw PM::TransformationError(msg);
    default: throw std::runtime_error(msg.empty() ? "b200reg: error " + std::to_string(status) : msg);
  }
}
static inline void check(b200reg_ctx* ctx, int32_t st) { if (st != B200REG_OK) throwStatus(ctx, st); }

namespace {
struct CloudGuard {   // frees a device cloud on scope exit (exceptions included)
  b200reg_ctx* ctx;
  b200reg_cloud* c = nullptr;
  explicit CloudGuard(b200reg_ctx* x) : ctx(x) {}
  ~CloudGuard() { if (c) b200reg_cloud_free(ctx, c); }
};

// DataPoints -> device cloud (features as they lie in memory; "normals" rows gathered into 3 floats per point)
void upload(b200reg_ctx* ctx, const PM::DataPoints& dp, bool withNormals, CloudGuard& g) {
  const int64_t n = dp.features.cols();
  std::vector<float> normals;
  const float* np = nullptr;
  if (withNormals && dp.descriptorExists("normals") && n > 0) {
    const int row = dp.getDescriptorStartingRow("normals");
    normals.resize(3 * (size_t)n);
    for (int64_t c = 0; c < n; ++c) for (int r = 0; r < 3; ++r) normals[3 * c + r] = dp.descriptors(row + r, c);
    np = normals.data();
  }
  check(ctx, b200reg_cloud_upload_f32(ctx, dp.features.data(), (int32_t)dp.features.rows(), np, n, &g.c));
}

std::vector<float> normalsOf(const PM::DataPoints& dp) {
  std::vector<float> out;
  if (!dp.descriptorExists("normals")) return out;
  const int64_t n = dp.features.cols();
  const int row = dp.getDescriptorStartingRow("normals");
  out.resize(3 * (size_t)n);
  for (int64_t c = 0; c < n; ++c) for (int r = 0; r < 3; ++r) out[3 * c + r] = dp.descriptors(row + r, c);
  return out;
}
}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// CudaVoxelGridDataPointsFilter
// ---------------------------------------------------------------------------------------------------------------
const PM::ParametersDoc CudaVoxelGridDataPointsFilter::availableParameters() {
  return {{"vSizeX", "Dimension of each voxel cell in x direction", "1.0", "0.001", "+inf", &P::Comp<float>},
          {"vSizeY", "Dimension of each voxel cell in y direction", "1.0", "0.001", "+inf", &P::Comp<float>},
          {"vSizeZ", "Dimension of each voxel cell in z direction", "1.0", "0.001", "+inf", &P::Comp<float>},
          {"useCentroid", "If 1 (true), down-sample by using centroid of voxel cell.  If false (0), use center of voxel cell.", "1", "0", "1", P::Comp<bool>},
          {"averageExistingDescriptors", "whether the filter keep the existing point descriptors and average them or should it drop them", "1", "0", "1", P::Comp<bool>},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaVoxelGridDataPointsFilter::CudaVoxelGridDataPointsFilter(const PM::Parameters& params)
    : PM::DataPointsFilter("CudaVoxelGridDataPointsFilter", availableParameters(), params), vSizeX(get<float>("vSizeX")), vSizeY(get<float>("vSizeY")),
      vSizeZ(get<float>("vSizeZ")), useCentroid(get<bool>("useCentroid")), averageExistingDescriptors(get<bool>("averageExistingDescriptors")),
      device(get<int>("device")) {}
PM::DataPoints CudaVoxelGridDataPointsFilter::filter(const PM::DataPoints& input) { PM::DataPoints o(input); inPlaceFilter(o); return o; }
void CudaVoxelGridDataPointsFilter::inPlaceFilter(PM::DataPoints& cloud) {
  if (cloud.features.rows() != 4) throw PM::InvalidParameter("CudaVoxelGridDataPointsFilter: only 3D clouds (4 x N features) are supported");
  // descriptors other than "normals" do not live on the device: refuse rather than silently drop them (VoxelGrid.cpp:81-89 validates too)
  size_t dsum = 0;
  for (auto& l : cloud.descriptorLabels) dsum += l.span;
  if ((int64_t)dsum != cloud.descriptors.rows()) throw PM::DataPoints::InvalidField("VoxelGridDataPointsFilter: Error, descriptor labels do not match descriptor data");
  for (auto& l : cloud.descriptorLabels)
    if (l.text != "normals") throw PM::DataPoints::InvalidField("CudaVoxelGridDataPointsFilter: descriptor '" + l.text + "' is not carried on the device (only 'normals')");
  if (cloud.times.rows() != 0) throw PM::DataPoints::InvalidField("CudaVoxelGridDataPointsFilter: times are not carried on the device");
  b200reg_ctx* ctx = contextFor(device);
  CloudGuard g(ctx);
  upload(ctx, cloud, true, g);
  check(ctx, b200reg_voxelgrid(ctx, g.c, vSizeX, vSizeY, vSizeZ, useCentroid, averageExistingDescriptors));
  const int64_t m = b200reg_cloud_size(g.c);
  const bool hasN = cloud.descriptorExists("normals");
  std::vector<float> nrm(hasN ? 3 * (size_t)m : 0);
  cloud.conservativeResize(m);
  check(ctx, b200reg_cloud_download_f32(ctx, g.c, cloud.features.data(), 4, hasN ? nrm.data() : nullptr));
  if (hasN) {
    const int row = cloud.getDescriptorStartingRow("normals");
    for (int64_t c = 0; c < m; ++c) for (int r = 0; r < 3; ++r) cloud.descriptors(row + r, c) = nrm[3 * c + r];
  }
}

// ---------------------------------------------------------------------------------------------------------------
// CudaSurfaceNormalDataPointsFilter
// ---------------------------------------------------------------------------------------------------------------
const PM::ParametersDoc CudaSurfaceNormalDataPointsFilter::availableParameters() {
  return {{"knn", "number of nearest neighbors to consider, including the point itself", "5", "3", "2147483647", &P::Comp<unsigned>},
          {"maxDist", "maximum distance to consider for neighbors", "inf", "0", "inf", &P::Comp<float>},
          {"epsilon", "approximation to use for the nearest-neighbor search", "0", "0", "inf", &P::Comp<float>},
          {"keepNormals", "whether the normals should be added as descriptors to the resulting cloud", "1"},
          {"keepDensities", "whether the point densities should be added as descriptors to the resulting cloud", "0"},
          {"keepEigenValues", "whether the eigen values should be added as descriptors to the resulting cloud", "0"},
          {"keepEigenVectors", "whether the eigen vectors should be added as descriptors to the resulting cloud", "0"},
          {"keepMatchedIds", "whether the identifiers of matches points should be added as descriptors to the resulting cloud", "0"},
          {"keepMeanDist", "whether the distance to the nearest neighbor mean should be added as descriptors to the resulting cloud", "0"},
          {"keepSurfaceDescriptors", "whether local surface descriptors (linearity, planarity and curvature) should be added to the resulting cloud", "1"},
          {"sortEigen", "whether the eigenvalues and eigenvectors should be sorted (ascending) based on the eigenvalues", "0"},
          {"smoothNormals", "whether the normal vector should be average with the nearest neighbors", "0"},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaSurfaceNormalDataPointsFilter::CudaSurfaceNormalDataPointsFilter(const PM::Parameters& params)
    : PM::DataPointsFilter("CudaSurfaceNormalDataPointsFilter", availableParameters(), params), knn(get<unsigned>("knn")), maxDist(get<float>("maxDist")),
      epsilon(get<float>("epsilon")), keepNormals(get<bool>("keepNormals")), keepDensities(get<bool>("keepDensities")),
      keepEigenValues(get<bool>("keepEigenValues")), keepEigenVectors(get<bool>("keepEigenVectors")), keepMatchedIds(get<bool>("keepMatchedIds")),
      keepMeanDist(get<bool>("keepMeanDist")), keepSurfaceDescriptors(get<bool>("keepSurfaceDescriptors")), sortEigen(get<bool>("sortEigen")),
      smoothNormals(get<bool>("smoothNormals")), device(get<int>("device")) {
  if (smoothNormals && !keepNormals) throw PM::InvalidParameter("CudaSurfaceNormalDataPointsFilter: smoothNormals needs keepNormals (the reference dereferences an empty optional, SurfaceNormal.cpp:303)");
}
PM::DataPoints CudaSurfaceNormalDataPointsFilter::filter(const PM::DataPoints& input) { PM::DataPoints o(input); inPlaceFilter(o); return o; }
void CudaSurfaceNormalDataPointsFilter::inPlaceFilter(PM::DataPoints& cloud) {
  if (cloud.features.rows() != 4) throw PM::InvalidParameter("CudaSurfaceNormalDataPointsFilter: only 3D clouds (4 x N features) are supported");
  size_t dsum = 0;
  for (auto& l : cloud.descriptorLabels) dsum += l.span;
  if ((int64_t)dsum != cloud.descriptors.rows()) throw PM::DataPoints::InvalidField("SurfaceNormalDataPointsFilter: Error, descriptor labels do not match descriptor data");
  const int64_t n = cloud.features.cols();
  // label order of the reference (SurfaceNormal.cpp:126-144); allocateDescriptor reuses / appends like DataPoints::allocateFields
  int rowN = -1, rowD = -1, rowE = -1, rowL = -1, rowP = -1, rowC = -1, rowV = -1, rowI = -1, rowM = -1;
  if (keepNormals) rowN = cloud.allocateDescriptor("normals", 3);
  if (keepDensities) rowD = cloud.allocateDescriptor("densities", 1);
  if (keepEigenValues) rowE = cloud.allocateDescriptor("eigValues", 3);
  if (keepEigenVectors) rowV = cloud.allocateDescriptor("eigVectors", 9);
  if (keepMatchedIds) rowI = cloud.allocateDescriptor("matchedIds", (int)knn);
  if (keepMeanDist) rowM = cloud.allocateDescriptor("meanDists", 1);
  if (keepSurfaceDescriptors) { rowL = cloud.allocateDescriptor("linearity", 1); rowP = cloud.allocateDescriptor("planarity", 1); rowC = cloud.allocateDescriptor("curvature", 1); }
  if (n == 0) return;
  b200reg_ctx* ctx = contextFor(device);
  CloudGuard g(ctx);
  check(ctx, b200reg_cloud_upload_f32(ctx, cloud.features.data(), 4, nullptr, n, &g.c));
  std::vector<float> dens(keepDensities ? (size_t)n : 0), eig(keepEigenValues ? 3 * (size_t)n : 0), lpc(keepSurfaceDescriptors ? 3 * (size_t)n : 0), nrm(3 * (size_t)n);
  std::vector<float> vec(keepEigenVectors ? 9 * (size_t)n : 0), mid(keepMatchedIds ? (size_t)knn * (size_t)n : 0), md(keepMeanDist ? (size_t)n : 0);
  int32_t degenerate = 0;
  b200reg_normals_options opt{};
  opt.knn = (int32_t)knn; opt.max_dist = maxDist; opt.epsilon = epsilon; opt.smooth_normals = smoothNormals ? 1 : 0;
  opt.densities = keepDensities ? dens.data() : nullptr;
  opt.eig_values = keepEigenValues ? eig.data() : nullptr;
  opt.eig_vectors = keepEigenVectors ? vec.data() : nullptr;
  opt.matched_ids = keepMatchedIds ? mid.data() : nullptr;
  opt.mean_dists = keepMeanDist ? md.data() : nullptr;
  opt.lin_plan_curv = keepSurfaceDescriptors ? lpc.data() : nullptr;
  check(ctx, b200reg_surface_normals_ex(ctx, g.c, &opt, &degenerate));
  check(ctx, b200reg_cloud_download_f32(ctx, g.c, nullptr, 4, nrm.data()));
  for (int64_t c = 0; c < n; ++c) {
    if (keepNormals) for (int r = 0; r < 3; ++r) cloud.descriptors(rowN + r, c) = nrm[3 * c + r];
    if (keepDensities) cloud.descriptors(rowD, c) = dens[c];
    if (keepEigenValues) for (int r = 0; r < 3; ++r) cloud.descriptors(rowE + r, c) = eig[3 * c + r];   // ascending (documented order; sortEigen is therefore always honoured)
    if (keepEigenVectors) for (int r = 0; r < 9; ++r) cloud.descriptors(rowV + r, c) = vec[9 * c + r];   // serializeEigVec order (utils.h:90-101)
    if (keepMatchedIds) for (unsigned r = 0; r < knn; ++r) cloud.descriptors(rowI + (int)r, c) = mid[(size_t)knn * c + r];
    if (keepMeanDist) cloud.descriptors(rowM, c) = md[c];
    if (keepSurfaceDescriptors) { cloud.descriptors(rowL, c) = lpc[3 * c]; cloud.descriptors(rowP, c) = lpc[3 * c + 1]; cloud.descriptors(rowC, c) = lpc[3 * c + 2]; }
  }
  (void)degenerate;
}

// ---------------------------------------------------------------------------------------------------------------
// CudaKDTreeMatcher
// ---------------------------------------------------------------------------------------------------------------
const PM::ParametersDoc CudaKDTreeMatcher::availableParameters() {
  return {{"knn", "number of nearest neighbors to consider it the reference", "1", "1", "2147483647", &P::Comp<unsigned>},
          {"epsilon", "approximation to use for the nearest-neighbor search (accepted, treated as 0: the search is exact)", "0", "0", "inf", &P::Comp<float>},
          {"searchType", "Nabo search type (accepted for YAML compatibility, ignored: the index is a radix-sorted voxel grid)", "1", "0", "2", &P::Comp<unsigned>},
          {"maxDist", "maximum distance to consider for neighbors", "inf", "0", "inf", &P::Comp<float>},
          {"cell", "voxel-grid cell size in metres; 0 = automatic", "0", "0", "inf", &P::Comp<float>},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaKDTreeMatcher::CudaKDTreeMatcher(const PM::Parameters& params)
    : PM::Matcher("CudaKDTreeMatcher", availableParameters(), params), knn((int)get<unsigned>("knn")), epsilon(get<float>("epsilon")),
      searchType(get<unsigned>("searchType")), maxDist(get<float>("maxDist")), cell(get<float>("cell")), device(get<int>("device")) {}
CudaKDTreeMatcher::~CudaKDTreeMatcher() {
  if (index) b200reg_index_free(contextFor(device), index);
}
void CudaKDTreeMatcher::init(const PM::DataPoints& filteredReference) {   // MatchersImpl.cpp:109-114
  b200reg_ctx* ctx = contextFor(device);
  if (index) { b200reg_index_free(ctx, index); index = nullptr; }
  CloudGuard g(ctx);
  upload(ctx, filteredReference, true, g);
  check(ctx, b200reg_index_build(ctx, g.c, cell, &index));   // the index keeps its own sorted copy; the cloud can go
}
PM::Matches CudaKDTreeMatcher::findClosests(const PM::DataPoints& filteredReading) {   // MatchersImpl.cpp:117-132
  if (!index) throw std::runtime_error("CudaKDTreeMatcher: init() must be called before findClosests()");
  b200reg_ctx* ctx = contextFor(device);
  const int64_t n = filteredReading.features.cols();
  PM::Matches matches(knn, n);
  CloudGuard g(ctx);
  upload(ctx, filteredReading, false, g);
  uint64_t visits = 0;
  check(ctx, b200reg_knn(ctx, index, g.c, nullptr, knn, maxDist, epsilon, 1 /*ALLOW_SELF_MATCH*/, matches.ids.data(), matches.dists.data(), &visits));
  this->visitCounter += (unsigned long)visits;
  return matches;
}

// ---------------------------------------------------------------------------------------------------------------
// CudaTrimmedDistOutlierFilter / CudaSurfaceNormalOutlierFilter
// ---------------------------------------------------------------------------------------------------------------
const PM::ParametersDoc CudaTrimmedDistOutlierFilter::availableParameters() {
  return {{"ratio", "percentage to keep", "0.85", "0.0000001", "1.0", &P::Comp<float>}, {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaTrimmedDistOutlierFilter::CudaTrimmedDistOutlierFilter(const PM::Parameters& params)
    : PM::OutlierFilter("CudaTrimmedDistOutlierFilter", availableParameters(), params), ratio(get<float>("ratio")), device(get<int>("device")) {}
PM::OutlierWeights CudaTrimmedDistOutlierFilter::compute(const PM::DataPoints&, const PM::DataPoints&, const PM::Matches& input) {
  b200reg_ctx* ctx = contextFor(device);
  PM::OutlierWeights w(input.dists.rows(), input.dists.cols());
  float tau = 0.f;
  check(ctx, b200reg_trimmed_dist(ctx, input.dists.data(), (int64_t)input.dists.d.size(), ratio, &tau, w.data()));
  return w;
}

CudaDistOutlierFilterBase::CudaDistOutlierFilterBase(const std::string& name, const PM::ParametersDoc& doc, const PM::Parameters& params, int kind_, const char* key)
    : PM::OutlierFilter(name, doc, params), kind(kind_), param(get<float>(key)), device(get<int>("device")) {}
PM::OutlierWeights CudaDistOutlierFilterBase::compute(const PM::DataPoints&, const PM::DataPoints&, const PM::Matches& input) {
  b200reg_ctx* ctx = contextFor(device);
  PM::OutlierWeights w(input.dists.rows(), input.dists.cols());
  check(ctx, b200reg_dist_outlier(ctx, kind, input.dists.data(), (int64_t)input.dists.d.size(), param, nullptr, w.data()));
  return w;
}
const PM::ParametersDoc CudaMaxDistOutlierFilter::availableParameters() {   // OutlierFiltersImpl.h:46-49
  return {{"maxDist", "threshold distance (Euclidean norm)", "1", "0.0000001", "inf", &P::Comp<float>}, {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaMaxDistOutlierFilter::CudaMaxDistOutlierFilter(const PM::Parameters& params)
    : CudaDistOutlierFilterBase("CudaMaxDistOutlierFilter", availableParameters(), params, B200REG_OUTLIER_MAX_DIST, "maxDist") {}
const PM::ParametersDoc CudaMinDistOutlierFilter::availableParameters() {   // OutlierFiltersImpl.h:64-67
  return {{"minDist", "threshold distance (Euclidean norm)", "1", "0.0000001", "inf", &P::Comp<float>}, {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaMinDistOutlierFilter::CudaMinDistOutlierFilter(const PM::Parameters& params)
    : CudaDistOutlierFilterBase("CudaMinDistOutlierFilter", availableParameters(), params, B200REG_OUTLIER_MIN_DIST, "minDist") {}
const PM::ParametersDoc CudaMedianDistOutlierFilter::availableParameters() {   // OutlierFiltersImpl.h:83-86
  return {{"factor", "points farther away factor * median will be considered outliers.", "3", "0.0000001", "inf", &P::Comp<float>},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaMedianDistOutlierFilter::CudaMedianDistOutlierFilter(const PM::Parameters& params)
    : CudaDistOutlierFilterBase("CudaMedianDistOutlierFilter", availableParameters(), params, B200REG_OUTLIER_MEDIAN_DIST, "factor") {}

const PM::ParametersDoc CudaSurfaceNormalOutlierFilter::availableParameters() {
  return {{"maxAngle", "Maximum authorised angle between the 2 surface normals (in radian)", "1.57", "0.0", "3.1416", &P::Comp<float>},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaSurfaceNormalOutlierFilter::CudaSurfaceNormalOutlierFilter(const PM::Parameters& params)
    : PM::OutlierFilter("CudaSurfaceNormalOutlierFilter", availableParameters(), params), maxAngle(get<float>("maxAngle")), device(get<int>("device")) {}
PM::OutlierWeights CudaSurfaceNormalOutlierFilter::compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference, const PM::Matches& input) {
  b200reg_ctx* ctx = contextFor(device);
  PM::OutlierWeights w(input.dists.rows(), input.dists.cols());
  const std::vector<float> nr = normalsOf(filteredReading), nf = normalsOf(filteredReference);
  check(ctx, b200reg_surface_normal_outlier(ctx, nr.empty() ? nullptr : nr.data(), nf.empty() ? nullptr : nf.data(), filteredReference.features.cols(),
                                            input.ids.data(), (int32_t)input.ids.rows(), input.ids.cols(), maxAngle, w.data()));
  return w;
}

// ---------------------------------------------------------------------------------------------------------------
// CudaPointToPlaneErrorMinimizer
// ---------------------------------------------------------------------------------------------------------------
const PM::ParametersDoc CudaPointToPlaneErrorMinimizer::availableParameters() {
  return {{"force2D", "If set to true(1), the minimization will be forced to give a solution in 2D (i.e., on the XY-plane) even with 3D inputs.", "0", "0", "1", &P::Comp<bool>},
          {"force4DOF", "If set to true(1), the minimization will optimize only yaw and translation, pitch and roll will follow the prior.", "0", "0", "1", &P::Comp<bool>},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaPointToPlaneErrorMinimizer::CudaPointToPlaneErrorMinimizer(const PM::Parameters& params)
    : PM::ErrorMinimizer("CudaPointToPlaneErrorMinimizer", availableParameters(), params), force2D(get<bool>("force2D")), force4DOF(get<bool>("force4DOF")),
      device(get<int>("device")) {}
CudaPointToPlaneErrorMinimizer::CudaPointToPlaneErrorMinimizer(const std::string& name, const PM::ParametersDoc& doc, const PM::Parameters& params)
    : PM::ErrorMinimizer(name, doc, params), force2D(get<bool>("force2D")), force4DOF(get<bool>("force4DOF")), device(get<int>("device")) {}
PM::TransformationParameters CudaPointToPlaneErrorMinimizer::compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference,
                                                                     const PM::OutlierWeights& outlierWeights, const PM::Matches& matches) {
  return computeOnDevice(filteredReading, filteredReference, outlierWeights, matches, 0.f, nullptr);
}
PM::TransformationParameters CudaPointToPlaneErrorMinimizer::computeOnDevice(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference,
                                                                             const PM::OutlierWeights& outlierWeights, const PM::Matches& matches,
                                                                             float sensorStdDev, PM::Matrix* covariance) {
  b200reg_ctx* ctx = contextFor(device);
  if (!filteredReference.descriptorExists("normals")) throw PM::DataPoints::InvalidField("Field normals not found");   // PointToPlane.cpp:115
  CloudGuard gr(ctx), gf(ctx);
  upload(ctx, filteredReading, false, gr);
  upload(ctx, filteredReference, true, gf);
  double A[36], b[6];
  b200reg_step_stats st;
  check(ctx, b200reg_point_to_plane(ctx, gr.c, gf.c, matches.ids.data(), matches.dists.data(), outlierWeights.data(), (int32_t)matches.ids.rows(),
                                    force2D ? 1 : 0, A, b, &st));
  this->pointUsedRatio = st.point_used_ratio;
  this->weightedPointUsedRatio = st.weighted_point_used_ratio;
  float T[16];
  b200reg_solve_point_to_plane(A, b, st.mean_reading, st.mean_reference, force2D ? 1 : 0, force4DOF ? 1 : 0, T, nullptr);
  PM::TransformationParameters out(4, 4);
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out(i, j) = T[4 * i + j];
  if (covariance) {   // PointToPlaneWithCov.cpp:62-66: estimateCovariance(mPts, out) on the centred ErrorElements
    double cov[36];
    check(ctx, b200reg_point_to_plane_covariance(ctx, gr.c, gf.c, matches.ids.data(), matches.dists.data(), outlierWeights.data(),
                                                 (int32_t)matches.ids.rows(), st.mean_reading, st.mean_reference, T, sensorStdDev, cov));
    *covariance = PM::Matrix(6, 6);
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) (*covariance)(i, j) = (float)cov[6 * i + j];
  }
  return out;
}

float CudaPointToPlaneErrorMinimizer::getResidualError(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference,
                                                       const PM::OutlierWeights& outlierWeights, const PM::Matches& matches) const {   // PointToPlane.cpp:410-422
  b200reg_ctx* ctx = contextFor(device);
  if (!filteredReference.descriptorExists("normals")) throw PM::DataPoints::InvalidField("Field normals not found");
  CloudGuard gr(ctx), gf(ctx);
  upload(ctx, filteredReading, false, gr);
  upload(ctx, filteredReference, true, gf);
  double res = 0.0;
  check(ctx, b200reg_point_to_plane_residual(ctx, gr.c, gf.c, matches.ids.data(), matches.dists.data(), outlierWeights.data(), (int32_t)matches.ids.rows(),
                                             force2D ? 1 : 0, &res, nullptr));
  return (float)res;
}

// ---------------------------------------------------------------------------------------------------------------
// CudaPointToPlaneWithCovErrorMinimizer, CudaPointToPointErrorMinimizer, CudaVarTrimmedDistOutlierFilter
// ---------------------------------------------------------------------------------------------------------------
const PM::ParametersDoc CudaPointToPlaneWithCovErrorMinimizer::availableParameters() {   // PointToPlaneWithCov.h:70-76 (+ the base's force4DOF)
  return {{"force2D", "If set to true(1), the minimization will be force to give a solution in 2D (i.e., on the XY-plane) even with 3D inputs.", "0", "0", "1", &P::Comp<bool>},
          {"force4DOF", "If set to true(1), the minimization will optimize only yaw and translation, pitch and roll will follow the prior.", "0", "0", "1", &P::Comp<bool>},
          {"sensorStdDev", "sensor standard deviation", "0.01", "0.", "inf", &P::Comp<float>},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaPointToPlaneWithCovErrorMinimizer::CudaPointToPlaneWithCovErrorMinimizer(const PM::Parameters& params)
    : CudaPointToPlaneErrorMinimizer("CudaPointToPlaneWithCovErrorMinimizer", availableParameters(), params), sensorStdDev(get<float>("sensorStdDev")) {}
PM::TransformationParameters CudaPointToPlaneWithCovErrorMinimizer::compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference,
                                                                            const PM::OutlierWeights& outlierWeights, const PM::Matches& matches) {
  return computeOnDevice(filteredReading, filteredReference, outlierWeights, matches, sensorStdDev, &covMatrix);
}

const PM::ParametersDoc CudaPointToPointErrorMinimizer::availableParameters() { return {{"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}}; }
CudaPointToPointErrorMinimizer::CudaPointToPointErrorMinimizer(const PM::Parameters& params)
    : PM::ErrorMinimizer("CudaPointToPointErrorMinimizer", availableParameters(), params), device(get<int>("device")) {}
PM::TransformationParameters CudaPointToPointErrorMinimizer::compute(const PM::DataPoints& filteredReading, const PM::DataPoints& filteredReference,
                                                                     const PM::OutlierWeights& outlierWeights, const PM::Matches& matches) {
  if (filteredReading.features.rows() != 4) throw PM::InvalidParameter("CudaPointToPointErrorMinimizer: only 3D clouds (4 x N features) are supported");
  b200reg_ctx* ctx = contextFor(device);
  CloudGuard gr(ctx), gf(ctx);
  upload(ctx, filteredReading, false, gr);
  upload(ctx, filteredReference, false, gf);
  float T[16];
  b200reg_step_stats st;
  check(ctx, b200reg_point_to_point(ctx, gr.c, gf.c, matches.ids.data(), matches.dists.data(), outlierWeights.data(), (int32_t)matches.ids.rows(), T, &st));
  this->pointUsedRatio = st.point_used_ratio;
  this->weightedPointUsedRatio = st.weighted_point_used_ratio;
  PM::TransformationParameters out(4, 4);
  for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out(i, j) = T[4 * i + j];
  return out;
}

const PM::ParametersDoc CudaVarTrimmedDistOutlierFilter::availableParameters() {   // OutlierFiltersImpl.h:153-159
  return {{"minRatio", "min ratio", "0.05", "0.0000001", "1", &P::Comp<float>},
          {"maxRatio", "max ratio", "0.99", "0.0000001", "1", &P::Comp<float>},
          {"lambda", "lambda (part of the term that balance the rmsd: 1/ratio^lambda", "2.35"},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaVarTrimmedDistOutlierFilter::CudaVarTrimmedDistOutlierFilter(const PM::Parameters& params)
    : PM::OutlierFilter("CudaVarTrimmedDistOutlierFilter", availableParameters(), params), minRatio(get<float>("minRatio")), maxRatio(get<float>("maxRatio")),
      lambda(get<float>("lambda")), device(get<int>("device")) {
  if (minRatio >= maxRatio)   // OutlierFiltersImpl.cpp:160-163
    throw PM::InvalidParameter("VarTrimmedDistOutlierFilter: minRatio (" + std::to_string(minRatio) + ") should be smaller than maxRatio (" + std::to_string(maxRatio) + ")");
}
PM::OutlierWeights CudaVarTrimmedDistOutlierFilter::compute(const PM::DataPoints&, const PM::DataPoints&, const PM::Matches& input) {
  b200reg_ctx* ctx = contextFor(device);
  PM::OutlierWeights w(input.dists.rows(), input.dists.cols());
  check(ctx, b200reg_var_trimmed_dist(ctx, input.dists.data(), (int64_t)input.dists.d.size(), minRatio, maxRatio, lambda, nullptr, nullptr, w.data()));
  return w;
}

// ---------------------------------------------------------------------------------------------------------------
// the compacting DataPointsFilters
// ---------------------------------------------------------------------------------------------------------------
CudaCompactingFilterBase::CudaCompactingFilterBase(const std::string& name, const PM::ParametersDoc& doc, const PM::Parameters& params)
    : PM::DataPointsFilter(name, doc, params), device(get<int>("device")) { std::memset(&fp, 0, sizeof(fp)); }
PM::DataPoints CudaCompactingFilterBase::filter(const PM::DataPoints& input) { PM::DataPoints o(input); inPlaceFilter(o); return o; }
void CudaCompactingFilterBase::inPlaceFilter(PM::DataPoints& cloud) {
  if (cloud.features.rows() != 4) throw PM::InvalidParameter(className + ": only 3D clouds (4 x N features) are supported");
  for (auto& l : cloud.descriptorLabels)
    if (l.text != "normals") throw PM::DataPoints::InvalidField(className + ": descriptor '" + l.text + "' is not carried on the device (only 'normals')");
  if (cloud.times.rows() != 0) throw PM::DataPoints::InvalidField(className + ": times are not carried on the device");
  b200reg_ctx* ctx = contextFor(device);
  CloudGuard g(ctx);
  upload(ctx, cloud, true, g);
  check(ctx, b200reg_cloud_filter(ctx, g.c, &fp));
  const int64_t m = b200reg_cloud_size(g.c);
  const bool hasN = cloud.descriptorExists("normals");
  std::vector<float> nrm(hasN ? 3 * (size_t)m : 0);
  cloud.conservativeResize(m);
  check(ctx, b200reg_cloud_download_f32(ctx, g.c, cloud.features.data(), 4, hasN ? nrm.data() : nullptr));
  if (hasN) {
    const int row = cloud.getDescriptorStartingRow("normals");
    for (int64_t c = 0; c < m; ++c) for (int r = 0; r < 3; ++r) cloud.descriptors(row + r, c) = nrm[3 * c + r];
  }
}
const PM::ParametersDoc CudaMaxDistDataPointsFilter::availableParameters() {   // MaxDist.h:57-62
  return {{"dim", "dimension on which the filter will be applied. x=0, y=1, z=2, radius=-1", "-1", "-1", "2", &P::Comp<int>},
          {"maxDist", "maximum distance authorized. If dim is set to -1 (radius), the absolute value of minDist will be used. All points beyond that will be filtered.", "1", "-inf", "inf", &P::Comp<float>},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaMaxDistDataPointsFilter::CudaMaxDistDataPointsFilter(const PM::Parameters& params)
    : CudaCompactingFilterBase("CudaMaxDistDataPointsFilter", availableParameters(), params) {
  fp.kind = B200REG_FILTER_MAX_DIST; fp.dim = get<int>("dim"); fp.dist = get<float>("maxDist");
}
const PM::ParametersDoc CudaMinDistDataPointsFilter::availableParameters() {   // MinDist.h:57-62
  return {{"dim", "dimension on which the filter will be applied. x=0, y=1, z=2, radius=-1", "-1", "-1", "2", &P::Comp<int>},
          {"minDist", "minimum value authorized. If dim is set to -1 (radius), the absolute value of minDist will be used. All points before that will be filtered.", "1", "-inf", "inf", &P::Comp<float>},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaMinDistDataPointsFilter::CudaMinDistDataPointsFilter(const PM::Parameters& params)
    : CudaCompactingFilterBase("CudaMinDistDataPointsFilter", availableParameters(), params) {
  fp.kind = B200REG_FILTER_MIN_DIST; fp.dim = get<int>("dim"); fp.dist = get<float>("minDist");
}
const PM::ParametersDoc CudaBoundingBoxDataPointsFilter::availableParameters() {   // BoundingBox.h:58-68
  return {{"xMin", "minimum value on x-axis defining one side of the bounding box", "-1", "-inf", "inf", &P::Comp<float>},
          {"xMax", "maximum value on x-axis defining one side of the bounding box", "1", "-inf", "inf", &P::Comp<float>},
          {"yMin", "minimum value on y-axis defining one side of the bounding box", "-1", "-inf", "inf", &P::Comp<float>},
          {"yMax", "maximum value on y-axis defining one side of the bounding box", "1", "-inf", "inf", &P::Comp<float>},
          {"zMin", "minimum value on z-axis defining one side of the bounding box", "-1", "-inf", "inf", &P::Comp<float>},
          {"zMax", "maximum value on z-axis defining one side of the bounding box", "1", "-inf", "inf", &P::Comp<float>},
          {"removeInside", "If set to true (1), remove points inside the bounding box; else (0), remove points outside the bounding box", "1", "0", "1", P::Comp<bool>},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaBoundingBoxDataPointsFilter::CudaBoundingBoxDataPointsFilter(const PM::Parameters& params)
    : CudaCompactingFilterBase("CudaBoundingBoxDataPointsFilter", availableParameters(), params) {
  fp.kind = B200REG_FILTER_BOUNDING_BOX;
  const char* keys[6] = {"xMin", "xMax", "yMin", "yMax", "zMin", "zMax"};
  for (int i = 0; i < 6; ++i) fp.box[i] = get<float>(keys[i]);
  fp.remove_inside = get<bool>("removeInside") ? 1 : 0;
}
const PM::ParametersDoc CudaRandomSamplingDataPointsFilter::availableParameters() {   // RandomSampling.h:59-64 (+ seed: the CPU twin draws it from std::random_device)
  return {{"prob", "Probability to keep a point, one over decimation factor ", "0.75", "0", "1", &P::Comp<double>},
          {"randomSamplingMethod", "Random sampling method: Direct RNG (0) (fastest), Uniform (1) (more accurate but slower)", "0", "0", "1", &P::Comp<int>},
          {"seed", "seed of the std::minstd_rand sequence the points are drawn against", "1", "0", "2147483646", &P::Comp<int64_t>},
          {"device", "CUDA device ordinal", "0", "0", "1023", &P::Comp<int>}};
}
CudaRandomSamplingDataPointsFilter::CudaRandomSamplingDataPointsFilter(const PM::Parameters& params)
    : CudaCompactingFilterBase("CudaRandomSamplingDataPointsFilter", availableParameters(), params) {
  fp.kind = B200REG_FILTER_RANDOM_SAMPLING;
  fp.prob = get<double>("prob");
  fp.random_method = get<int>("randomSamplingMethod");
  fp.seed = (uint64_t)get<int64_t>("seed");
}

// ---------------------------------------------------------------------------------------------------------------
// registration (the ADD_TO_REGISTRAR lines a maintainer adds to Registry.cpp, or calls out-of-tree — INTEGRATION.md)
// ---------------------------------------------------------------------------------------------------------------
void registerCudaModules(PM& pm) {
  using PointMatcherSupport::Registrar;
  auto& DataPointsFilterRegistrar = pm.DataPointsFilterRegistrar;
  auto& MatcherRegistrar = pm.MatcherRegistrar;
  auto& OutlierFilterRegistrar = pm.OutlierFilterRegistrar;
  auto& ErrorMinimizerRegistrar = pm.ErrorMinimizerRegistrar;
  typedef PM::DataPointsFilter DataPointsFilter;
  typedef PM::Matcher Matcher;
  typedef PM::OutlierFilter OutlierFilter;
  typedef PM::ErrorMinimizer ErrorMinimizer;
  ADD_TO_REGISTRAR(DataPointsFilter, CudaVoxelGridDataPointsFilter, CudaVoxelGridDataPointsFilter)
  ADD_TO_REGISTRAR(DataPointsFilter, CudaSurfaceNormalDataPointsFilter, CudaSurfaceNormalDataPointsFilter)
  ADD_TO_REGISTRAR(Matcher, CudaKDTreeMatcher, CudaKDTreeMatcher)
  ADD_TO_REGISTRAR(OutlierFilter, CudaTrimmedDistOutlierFilter, CudaTrimmedDistOutlierFilter)
  ADD_TO_REGISTRAR(OutlierFilter, CudaSurfaceNormalOutlierFilter, CudaSurfaceNormalOutlierFilter)
  ADD_TO_REGISTRAR(OutlierFilter, CudaMaxDistOutlierFilter, CudaMaxDistOutlierFilter)
  ADD_TO_REGISTRAR(OutlierFilter, CudaMinDistOutlierFilter, CudaMinDistOutlierFilter)
  ADD_TO_REGISTRAR(OutlierFilter, CudaMedianDistOutlierFilter, CudaMedianDistOutlierFilter)
  ADD_TO_REGISTRAR(ErrorMinimizer, CudaPointToPlaneErrorMinimizer, CudaPointToPlaneErrorMinimizer)
  ADD_TO_REGISTRAR(ErrorMinimizer, CudaPointToPlaneWithCovErrorMinimizer, CudaPointToPlaneWithCovErrorMinimizer)
  ADD_TO_REGISTRAR(ErrorMinimizer, CudaPointToPointErrorMinimizer, CudaPointToPointErrorMinimizer)
  ADD_TO_REGISTRAR(OutlierFilter, CudaVarTrimmedDistOutlierFilter, CudaVarTrimmedDistOutlierFilter)
  ADD_TO_REGISTRAR(DataPointsFilter, CudaMaxDistDataPointsFilter, CudaMaxDistDataPointsFilter)
  ADD_TO_REGISTRAR(DataPointsFilter, CudaMinDistDataPointsFilter, CudaMinDistDataPointsFilter)
  ADD_TO_REGISTRAR(DataPointsFilter, CudaBoundingBoxDataPointsFilter, CudaBoundingBoxDataPointsFilter)
  ADD_TO_REGISTRAR(DataPointsFilter, CudaRandomSamplingDataPointsFilter, CudaRandomSamplingDataPointsFilter)
}

}  // namespace b200
