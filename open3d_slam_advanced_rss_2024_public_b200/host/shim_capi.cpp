// host/shim_capi.cpp — a small extern "C" window onto the C++ shim so that pytest (ctypes) can drive the registrable
// modules exactly as a libpointmatcher user would: build a PM::ICP from YAML text, run it on DataPoints, catch the
// libpointmatcher exception types.  Test/harness glue only; product users include cuda_modules.h directly.
#include <cstring>
#include <sstream>

#include "cuda_modules.h"

namespace {

enum { EXC_NONE = 0, EXC_INVALID_PARAMETER = 1, EXC_INVALID_ELEMENT = 2, EXC_INVALID_MODULE_TYPE = 3, EXC_CONVERGENCE = 4,
       EXC_TRANSFORMATION = 5, EXC_INVALID_FIELD = 6, EXC_CONFIGURATION = 7, EXC_RUNTIME = 8 };

void setMsg(char* buf, int cap, const std::string& s) {
  if (!buf || cap <= 0) return;
  std::strncpy(buf, s.c_str(), (size_t)cap - 1);
  buf[cap - 1] = 0;
}

template <typename F>
int guarded(char* err, int cap, F&& f) {
  try {
    f();
    setMsg(err, cap, "");
    return EXC_NONE;
  } catch (const PointMatcherSupport::InvalidParameter& e) { setMsg(err, cap, e.what()); return EXC_INVALID_PARAMETER;
  } catch (const PointMatcherSupport::InvalidElement& e) { setMsg(err, cap, e.what()); return EXC_INVALID_ELEMENT;
  } catch (const PM::InvalidModuleType& e) { setMsg(err, cap, e.what()); return EXC_INVALID_MODULE_TYPE;
  } catch (const PM::ConvergenceError& e) { setMsg(err, cap, e.what()); return EXC_CONVERGENCE;
  } catch (const PM::TransformationError& e) { setMsg(err, cap, e.what()); return EXC_TRANSFORMATION;
  } catch (const PM::DataPoints::InvalidField& e) { setMsg(err, cap, e.what()); return EXC_INVALID_FIELD;
  } catch (const PointMatcherSupport::ConfigurationError& e) { setMsg(err, cap, e.what()); return EXC_CONFIGURATION;
  } catch (const std::runtime_error& e) { setMsg(err, cap, e.what()); return EXC_RUNTIME; }
}

// N x stride floats (xyz first) + optional N x 3 normals -> DataPoints with 4 x N features and a "normals" descriptor
PM::DataPoints makeDataPoints(const float* feat, int stride, const float* normals, int64_t n) {
  PM::DataPoints dp;
  dp.features = PM::Matrix(4, n);
  dp.featureLabels = {PM::DataPoints::Label("x", 1), PM::DataPoints::Label("y", 1), PM::DataPoints::Label("z", 1), PM::DataPoints::Label("pad", 1)};
  for (int64_t c = 0; c < n; ++c) {
    for (int r = 0; r < 3; ++r) dp.features(r, c) = feat[c * stride + r];
    dp.features(3, c) = 1.f;
  }
  if (normals) {
    dp.descriptors = PM::Matrix(3, n);
    dp.descriptorLabels = {PM::DataPoints::Label("normals", 3)};
    for (int64_t c = 0; c < n; ++c) for (int r = 0; r < 3; ++r) dp.descriptors(r, c) = normals[3 * c + r];
  }
  return dp;
}

struct IcpHandle {
  PM::ICP icp;
};

}  // namespace

extern "C" {

// ---- registrar introspection ----
// kinds: 0 DataPointsFilter, 1 Matcher, 2 OutlierFilter, 3 ErrorMinimizer, 4 TransformationChecker, 5 Transformation, 6 Inspector, 7 Logger
int b200pm_list_modules(int kind, char* out, int cap) {
  const PM& pm = PM::get();
  std::vector<std::string> names;
  switch (kind) {
    case 0: names = pm.DataPointsFilterRegistrar.names(); break;
    case 1: names = pm.MatcherRegistrar.names(); break;
    case 2: names = pm.OutlierFilterRegistrar.names(); break;
    case 3: names = pm.ErrorMinimizerRegistrar.names(); break;
    case 4: names = pm.TransformationCheckerRegistrar.names(); break;
    case 5: names = pm.TransformationRegistrar.names(); break;
    case 6: names = pm.InspectorRegistrar.names(); break;
    case 7: names = pm.LoggerRegistrar.names(); break;
    default: return -1;
  }
  std::string s;
  for (auto& n : names) s += n + "\n";
  setMsg(out, cap, s);
  return (int)names.size();
}

// parameter table of a module ("name|default|min|max\n" per parameter) — the static contract of SURVEY 8b
int b200pm_module_parameters(int kind, const char* name, char* out, int cap, char* err, int errcap) {
  return guarded(err, errcap, [&] {
    const PM& pm = PM::get();
    PM::ParametersDoc doc;
    switch (kind) {
      case 0: doc = pm.DataPointsFilterRegistrar.getDescriptor(name)->availableParameters(); break;
      case 1: doc = pm.MatcherRegistrar.getDescriptor(name)->availableParameters(); break;
      case 2: doc = pm.OutlierFilterRegistrar.getDescriptor(name)->availableParameters(); break;
      case 3: doc = pm.ErrorMinimizerRegistrar.getDescriptor(name)->availableParameters(); break;
      case 4: doc = pm.TransformationCheckerRegistrar.getDescriptor(name)->availableParameters(); break;
      default: throw PointMatcherSupport::InvalidElement("unknown registrar kind");
    }
    std::string s;
    for (auto& p : doc) s += p.name + "|" + p.defaultValue + "|" + p.minValue + "|" + p.maxValue + "\n";
    setMsg(out, cap, s);
  });
}

// ---- ICP from YAML ----
void* b200pm_icp_create() { return new IcpHandle(); }
void b200pm_icp_destroy(void* h) { delete (IcpHandle*)h; }

int b200pm_icp_load_yaml(void* h, const char* yaml, char* err, int errcap) {
  return guarded(err, errcap, [&] {
    std::istringstream is(yaml);
    ((IcpHandle*)h)->icp.loadFromYaml(is);
  });
}
int b200pm_icp_set_default(void* h, char* err, int errcap) {
  return guarded(err, errcap, [&] { ((IcpHandle*)h)->icp.setDefault(); });
}

// summary of the loaded chain: "section:ClassName" lines (what loadFromYaml instantiated through the registrars)
int b200pm_icp_describe(void* h, char* out, int cap) {
  PM::ICP& icp = ((IcpHandle*)h)->icp;
  std::string s;
  for (auto& f : icp.readingDataPointsFilters) s += "readingDataPointsFilters:" + f->className + "\n";
  for (auto& f : icp.readingStepDataPointsFilters) s += "readingStepDataPointsFilters:" + f->className + "\n";
  for (auto& f : icp.referenceDataPointsFilters) s += "referenceDataPointsFilters:" + f->className + "\n";
  if (icp.matcher) s += "matcher:" + icp.matcher->className + "\n";
  for (auto& f : icp.outlierFilters) s += "outlierFilters:" + f->className + "\n";
  if (icp.errorMinimizer) s += "errorMinimizer:" + icp.errorMinimizer->className + "\n";
  for (auto& f : icp.transformationCheckers) s += "transformationCheckers:" + f->className + "\n";
  setMsg(out, cap, s);
  return 0;
}

// T = icp(reading, reference, T_init)   (examples/icp.cpp:146 / Mapper.cpp:391-393 usage).  4x4 matrices row-major.
// force_generic != 0 swaps the matcher's parameters?  No: it disables the fast lane by inserting an IdentityDataPointsFilter
// into readingStepDataPointsFilters, so the same chain runs module by module through the Eigen-shaped interfaces.
int b200pm_icp_run(void* h, const float* ref_feat, int ref_stride, const float* ref_normals, int64_t m, const float* read_feat,
                   int read_stride, const float* read_normals, int64_t n, const float* T_init, int force_generic, float* T_out,
                   int* iterations, int* used_fast_lane, int* max_iter_reached, float* overlap, char* err, int errcap) {
  return guarded(err, errcap, [&] {
    PM::ICP& icp = ((IcpHandle*)h)->icp;
    const PM::DataPoints ref = makeDataPoints(ref_feat, ref_stride, ref_normals, m);
    const PM::DataPoints reading = makeDataPoints(read_feat, read_stride, read_normals, n);
    PM::TransformationParameters Ti = PM::Matrix::Identity(4);
    if (T_init) for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) Ti(i, j) = T_init[4 * i + j];
    const size_t nStep = icp.readingStepDataPointsFilters.size();
    if (force_generic) icp.readingStepDataPointsFilters.push_back(PM::get().DataPointsFilterRegistrar.create("IdentityDataPointsFilter"));
    struct Restore { PM::ICP& i; size_t n; ~Restore() { i.readingStepDataPointsFilters.resize(n); } } restore{icp, nStep};
    const PM::TransformationParameters T = icp(reading, ref, Ti);
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) T_out[4 * i + j] = T(i, j);
    if (iterations) *iterations = icp.getIterationCount();
    if (used_fast_lane) *used_fast_lane = icp.usedFastLane() ? 1 : 0;
    if (max_iter_reached) *max_iter_reached = icp.getMaxNumIterationsReached() ? 1 : 0;
    if (overlap) *overlap = icp.errorMinimizer->getWeightedPointUsedRatio();
  });
}

// ---- a single DataPointsFilter by name (swap-one-module test shape of the reference's utest) ----
// in/out: feat N x 4, normals N x 3 (has_normals says whether the input carries them); returns the new N via n_out;
// extra descriptors produced by the filter are returned by name through `desc_names` ("name:span\n").
int b200pm_apply_filter(const char* name, const char* params_kv /* "k=v\n..." */, float* feat, float* normals, int has_normals, int64_t n,
                        int64_t* n_out, int* normals_out, char* desc_names, int cap, char* err, int errcap) {
  return guarded(err, errcap, [&] {
    PM::Parameters params;
    std::istringstream is(params_kv ? params_kv : "");
    std::string line;
    while (std::getline(is, line)) {
      const size_t eq = line.find('=');
      if (eq != std::string::npos) params[line.substr(0, eq)] = line.substr(eq + 1);
    }
    auto filter = PM::get().DataPointsFilterRegistrar.create(name, params);
    PM::DataPoints dp = makeDataPoints(feat, 4, has_normals ? normals : nullptr, n);
    filter->init();
    filter->inPlaceFilter(dp);
    const int64_t m = dp.features.cols();
    *n_out = m;
    for (int64_t c = 0; c < m; ++c) for (int r = 0; r < 4; ++r) feat[4 * c + r] = dp.features(r, c);
    *normals_out = dp.descriptorExists("normals") ? 1 : 0;
    if (*normals_out) {
      const int row = dp.getDescriptorStartingRow("normals");
      for (int64_t c = 0; c < m; ++c) for (int r = 0; r < 3; ++r) normals[3 * c + r] = dp.descriptors(row + r, c);
    }
    std::string s;
    for (auto& l : dp.descriptorLabels) s += l.text + ":" + std::to_string(l.span) + "\n";
    setMsg(desc_names, cap, s);
  });
}

}  // extern "C"

// ---- open3d_slam Mapper driver (host/mapper.h) ----------------------------------------------------------------------
#include "mapper.h"

extern "C" {

typedef struct b200map_params {   // flat mirror of b200::MapperParams for ctypes
  b200reg_crop_spec map_builder_cropper;
  b200reg_crop_spec scan_matcher_cropper;
  double scan_voxel_size, normals_radius, map_voxel_size, min_movement, reference_period;
  int32_t normals_knn, pad;
  b200reg_icp_config icp;
  int32_t carving_enabled, carve_every;
  b200reg_carve_params carving;
} b200map_params;

typedef struct b200map_stats {
  int32_t iterations, icp_status, reference_refreshed, inserted;
  int64_t reading_points, merge_points, patch_points, map_points;
  double stage_ms[9];   // MapperStepStats::stageMs
  int64_t carved_points;
  double prefetch_wait_ms;
  int64_t prefetched;
} b200map_stats;

void* b200map_create(b200reg_ctx* ctx, const b200map_params* p) {
  if (!ctx || !p) return nullptr;
  b200::MapperParams mp;
  mp.mapBuilderCropper = p->map_builder_cropper;
  mp.scanMatcherCropper = p->scan_matcher_cropper;
  mp.scanVoxelSize = p->scan_voxel_size;
  mp.normalsKnn = p->normals_knn;
  mp.normalsRadius = p->normals_radius;
  mp.mapVoxelSize = p->map_voxel_size;
  mp.minMovementBetweenMappingSteps = p->min_movement;
  mp.referenceCloudSettingPeriod = p->reference_period;
  mp.icp = p->icp;
  mp.carvingEnabled = p->carving_enabled != 0;
  mp.carveSpaceEveryNscans = p->carve_every;
  mp.carving = p->carving;
  return new b200::Mapper(ctx, mp);
}
void b200map_destroy(void* m) { delete (b200::Mapper*)m; }
void b200map_set_pose(void* m, const double T[16]) { ((b200::Mapper*)m)->setMapToRangeSensor(T); }
int32_t b200map_add_range_measurement(void* m, const double* xyz, const double* normals, int64_t n, double time, const double odom[16],
                                      double pose_out[16], b200map_stats* stats, char* err, int32_t err_cap) {
  b200::Mapper* mp = (b200::Mapper*)m;
  const bool ok = mp->addRangeMeasurement(xyz, normals, n, time, odom);
  std::memcpy(pose_out, mp->mapToRangeSensor(), sizeof(double) * 16);
  if (stats) {
    const b200::MapperStepStats& s = mp->lastStats();
    *stats = b200map_stats{s.iterations, s.icpStatus, s.referenceRefreshed, s.inserted, s.readingPoints, s.mergePoints, s.patchPoints, s.mapPoints, {}};
    for (int i = 0; i < b200::MapperStepStats::kStages; ++i) stats->stage_ms[i] = s.stageMs[i];
    stats->carved_points = s.carvedPoints;
    stats->prefetch_wait_ms = s.prefetchWaitMs;
    stats->prefetched = s.prefetched;
  }
  setMsg(err, err_cap, mp->lastError());
  return ok ? 1 : 0;
}
int32_t b200map_add_pointcloud2(void* m, const uint8_t* data, size_t bytes, const b200reg_pc2_layout* layout, double time, const double odom[16],
                                double pose_out[16], b200map_stats* stats, char* err, int32_t err_cap) {
  b200::Mapper* mp = (b200::Mapper*)m;
  const bool ok = layout && mp->addRangeMeasurementPointCloud2(data, bytes, *layout, time, odom);
  std::memcpy(pose_out, mp->mapToRangeSensor(), sizeof(double) * 16);
  if (stats) {
    const b200::MapperStepStats& s = mp->lastStats();
    *stats = b200map_stats{s.iterations, s.icpStatus, s.referenceRefreshed, s.inserted, s.readingPoints, s.mergePoints, s.patchPoints, s.mapPoints, {}};
    for (int i = 0; i < b200::MapperStepStats::kStages; ++i) stats->stage_ms[i] = s.stageMs[i];
    stats->carved_points = s.carvedPoints;
    stats->prefetch_wait_ms = s.prefetchWaitMs;
    stats->prefetched = s.prefetched;
  }
  setMsg(err, err_cap, mp->lastError());
  return ok ? 1 : 0;
}
int32_t b200map_prefetch(void* m, const double* xyz, int64_t n) { return ((b200::Mapper*)m)->prefetch(xyz, n) ? 1 : 0; }
b200reg_map* b200map_map(void* m) { return ((b200::Mapper*)m)->map(); }

}  // extern "C"
