// host/mapper.cpp — see mapper.h.  Poses are fp64 4x4 row-major on this side and are cast to fp32 for ICP and back
// (Mapper.cpp:323,435).
#include "mapper.h"

#include <chrono>
#include <cmath>
#include <cstring>
#include <utility>

namespace b200 {
namespace {

void identity(double* T) {
  for (int i = 0; i < 16; ++i) T[i] = (i % 5 == 0) ? 1.0 : 0.0;
}
void mul(const double* a, const double* b, double* r) {
  double t[16];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      double acc = 0.0;
      for (int k = 0; k < 4; ++k) acc += a[4 * i + k] * b[4 * k + j];
      t[4 * i + j] = acc;
    }
  std::memcpy(r, t, sizeof(t));
}
// Eigen::Transform<double,3,Affine>::inverse(): linear part by the closed-form 3x3 inverse, translation = -inv(L) t
void inverseAffine(const double* T, double* r) {
  const double a = T[0], b = T[1], c = T[2], d = T[4], e = T[5], f = T[6], g = T[8], h = T[9], k = T[10];
  const double c00 = e * k - f * h, c01 = c * h - b * k, c02 = b * f - c * e;
  const double c10 = f * g - d * k, c11 = a * k - c * g, c12 = c * d - a * f;
  const double c20 = d * h - e * g, c21 = b * g - a * h, c22 = a * e - b * d;
  const double det = a * c00 + b * c10 + c * c20, inv = 1.0 / det;
  const double L[9] = {c00 * inv, c01 * inv, c02 * inv, c10 * inv, c11 * inv, c12 * inv, c20 * inv, c21 * inv, c22 * inv};
  identity(r);
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) r[4 * i + j] = L[3 * i + j];
    r[4 * i + 3] = -(L[3 * i] * T[3] + L[3 * i + 1] * T[7] + L[3 * i + 2] * T[11]);
  }
}

b200reg_crop_spec at(const b200reg_crop_spec& s, const double* T) {   // CroppingVolume::setPose: only the translation matters
  b200reg_crop_spec o = s;
  o.center[0] = T ? T[3] : 0.0;
  o.center[1] = T ? T[7] : 0.0;
  o.center[2] = T ? T[11] : 0.0;
  return o;
}

struct StageClock {   // adds the wall time since construction / the last lap to one slot of MapperStepStats::stageMs
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  void lap(double* slot) {
    const auto t1 = std::chrono::steady_clock::now();
    *slot += std::chrono::duration<double, std::milli>(t1 - t0).count();
    t0 = t1;
  }
};

}  // namespace

Mapper::Mapper(b200reg_ctx* ctx, const MapperParams& params) : ctx_(ctx), params_(params) {
  identity(mapToRangeSensor_);
  identity(mapToRangeSensorPrev_);
  identity(lastInsertion_);
  identity(odomPrev_);
  identity(cropperPose_);
  if (b200reg_map_create(ctx_, params_.mapVoxelSize, &map_) != B200REG_OK) error_ = b200reg_last_error(ctx_);
  if (b200reg_icp_create(ctx_, &params_.icp, &icp_) != B200REG_OK) error_ = b200reg_last_error(ctx_);
}

void Mapper::freeSet(ScanSet& s) {
  if (s.reading) b200reg_cloud_free(ctx_, s.reading);
  b200reg_cloud64_free(ctx_, s.raw);
  b200reg_cloud64_free(ctx_, s.cropped);
  b200reg_cloud64_free(ctx_, s.merge);
  b200reg_cloud64_free(ctx_, s.match);
  s = ScanSet();
}

Mapper::~Mapper() {
  if (workersUp_) {
    { std::lock_guard<std::mutex> lock(mu_); stop_ = true; }
    cv_.notify_all();
    for (std::thread& t : worker_) if (t.joinable()) t.join();
  }
  for (b200reg_ctx* c : workerCtx_) if (c) b200reg_ctx_sync(c);
  for (ScanSet& s : sets_) freeSet(s);
  if (icp_) b200reg_icp_destroy(icp_);
  if (map_) b200reg_map_destroy(ctx_, map_);
  for (b200reg_ctx* c : workerCtx_) if (c) b200reg_ctx_destroy(c);
}

void Mapper::setMapToRangeSensor(const double T[16]) {
  std::memcpy(mapToRangeSensor_, T, sizeof(mapToRangeSensor_));
  std::memcpy(mapToRangeSensorPrev_, T, sizeof(mapToRangeSensorPrev_));
}

// ScanToMapIcp::processForScanMatchingAndMerging (ScanToMapRegistration.cpp:36-69), every step on the device: the raw scan is
// the only host -> device copy of a mapping step
bool Mapper::process(b200reg_ctx* ctx, ScanSet& s) {
  StageClock clk;
  auto fail = [&]() { s.error = b200reg_last_error(ctx); return false; };
  auto say = [&](const char* msg) { s.error = msg; return false; };
  // wide crop: mapBuilderCropper_->crop(in) — that cropper instance never gets a pose, i.e. it sits at the identity (:30-37)
  const b200reg_crop_spec wide = at(params_.mapBuilderCropper, nullptr);
  if (b200reg_cloud64_crop(ctx, &wide, s.raw, &s.cropped) != B200REG_OK) return fail();
  clk.lap(&s.stageMs[0]);
  // o3d_slam::voxelize (helpers.cpp:108-115): no-op for voxel <= 0
  if (params_.scanVoxelSize > 0.0 && b200reg_cloud64_size(s.cropped) > 0) {
    if (b200reg_cloud64_voxel_down_sample(ctx, params_.scanVoxelSize, s.cropped, &s.merge) != B200REG_OK) return fail();
  } else {
    std::swap(s.cropped, s.merge);
    if (!s.merge) return say("ScanToMapIcp::wideCropped cropped size is zero");
  }
  if (b200reg_cloud64_size(s.merge) == 0) return say("ScanToMapIcp::wideCropped cropped size is zero");   // assert_gt (:66)
  clk.lap(&s.stageMs[1]);
  // estimateNormalsOrCovariancesIfNeeded (CloudRegistration.cpp:62-75): only when the cloud has no normals;
  // EstimateNormals(Hybrid(radius, knn)) + NormalizeNormals + OrientNormalsTowardsCameraLocation(origin)
  if (!b200reg_cloud64_has_normals(s.merge)) {
    const double origin[3] = {0.0, 0.0, 0.0};
    int32_t degenerate = 0;
    if (b200reg_cloud64_estimate_normals(ctx, s.merge, params_.normalsKnn, params_.normalsRadius, origin, &degenerate) != B200REG_OK) return fail();
  }
  clk.lap(&s.stageMs[2]);
  // narrow crop at the identity pose (:62-63)
  const b200reg_crop_spec narrow = at(params_.scanMatcherCropper, nullptr);
  if (b200reg_cloud64_crop(ctx, &narrow, s.merge, &s.match) != B200REG_OK) return fail();
  if (b200reg_cloud64_size(s.match) == 0) return say("ScanToMapIcp::narrow cropped size is zero");   // assert_gt (:65)
  clk.lap(&s.stageMs[3]);
  // open3dToPointmatcher (Mapper.cpp:307-309), device to device
  if (b200reg_cloud64_to_f32(ctx, s.match, &s.reading) != B200REG_OK) return fail();
  clk.lap(&s.stageMs[4]);
  return true;
}

// A worker of prefetch(): takes the next queued set, uploads the scan and runs the chain on its own context, marks the set ready.
void Mapper::workerLoop(int w) {
  b200reg_ctx* ctx = workerCtx_[w];
  for (;;) {
    int j;
    {
      std::unique_lock<std::mutex> lock(mu_);
      cv_.wait(lock, [&] { return stop_ || !jobs_.empty(); });
      if (stop_) return;
      j = jobs_.front();
      jobs_.pop_front();
    }
    ScanSet& s = sets_[j];
    StageClock clk;
    for (double& v : s.stageMs) v = 0.0;
    s.error.clear();
    bool ok = b200reg_cloud64_upload(ctx, s.srcPtr, nullptr, s.srcN, &s.raw) == B200REG_OK;
    if (!ok) s.error = b200reg_last_error(ctx);
    clk.lap(&s.stageMs[0]);   // (the upload is booked with the first stage)
    if (ok) ok = process(ctx, s);
    if (b200reg_ctx_sync(ctx) != B200REG_OK && ok) { ok = false; s.error = b200reg_last_error(ctx); }
    {
      std::lock_guard<std::mutex> lock(mu_);
      s.ok = ok;
      s.state = 2;
    }
    cv_.notify_all();
  }
}

// a set nobody is waiting for: a free one, else the oldest pending one once its worker has finished with it (mu_ held)
int Mapper::takeSetLocked(std::unique_lock<std::mutex>& lock) {
  for (int k = 0; k < kSets; ++k)
    if (sets_[k].state == 0) return k;
  int j = 0;
  for (int k = 1; k < kSets; ++k)
    if (sets_[k].ticket < sets_[j].ticket) j = k;
  cv_.wait(lock, [&] { return sets_[j].state == 2; });
  sets_[j].state = 0;
  sets_[j].srcPtr = nullptr;
  return j;
}

bool Mapper::prefetch(const double* xyz, int64_t n) {
  if (!xyz || n < 0) { error_ = "prefetch: bad arguments"; return false; }
  if (!workersUp_) {
    for (int w = 0; w < kWorkers; ++w)
      if (b200reg_ctx_create(b200reg_ctx_device(ctx_), &workerCtx_[w]) != B200REG_OK) {
        error_ = "prefetch: cannot create a worker context";
        for (b200reg_ctx*& c : workerCtx_) { if (c) b200reg_ctx_destroy(c); c = nullptr; }
        return false;
      }
    for (int w = 0; w < kWorkers; ++w) worker_[w] = std::thread(&Mapper::workerLoop, this, w);
    workersUp_ = true;
  }
  {
    std::unique_lock<std::mutex> lock(mu_);
    const int j = takeSetLocked(lock);
    ScanSet& s = sets_[j];
    s.srcPtr = xyz;
    s.srcN = n;
    s.state = 1;
    s.ticket = ++tickets_;
    jobs_.push_back(j);
  }
  cv_.notify_all();
  return true;
}

// Submap::insertScan (Submap.cpp:39-96) without carving: transform, append, re-voxelise inside the map-builder volume at the pose
bool Mapper::insert() {
  StageClock clk;
  // carve(rawScan, mapToRangeSensor, *mapBuilderCropper_, carving_, &mapCloud_) (Submap.cpp:56-60,130-144): every N-th inserted scan,
  // with the cropper still at the pose of the previous insertion
  if (params_.carvingEnabled && b200reg_map_size(map_) > 0 && params_.carveSpaceEveryNscans > 0 &&
      nScansInsertedMap_ % params_.carveSpaceEveryNscans == 0) {
    const b200reg_crop_spec carveVol = at(params_.mapBuilderCropper, cropperPose_);
    if (b200reg_map_carve_cloud64(ctx_, map_, sets_[cur_].raw, mapToRangeSensor_, &carveVol, &params_.carving, &stats_.carvedPoints) != B200REG_OK) {
      error_ = b200reg_last_error(ctx_);
      return false;
    }
  }
  const b200reg_crop_spec vol = at(params_.mapBuilderCropper, mapToRangeSensor_);
  if (b200reg_map_insert_cloud64(ctx_, map_, sets_[cur_].merge, mapToRangeSensor_, &vol) != B200REG_OK) {
    const std::string earlier = error_;   // a failed registration only keeps the estimate; its message must not be lost here
    error_ = b200reg_last_error(ctx_);
    if (!earlier.empty()) error_ = earlier + " | " + error_;
    return false;
  }
  std::memcpy(cropperPose_, mapToRangeSensor_, sizeof(cropperPose_));   // mapBuilderCropper_->setPose(mapToRangeSensor) (Submap.cpp:86)
  ++nScansInsertedMap_;
  clk.lap(&stats_.stageMs[MapperStepStats::kInsert]);
  return true;
}

bool Mapper::addRangeMeasurement(const double* xyz, const double* normals, int64_t n, double time, const double odom[16]) {
  stats_ = MapperStepStats();
  error_.clear();
  StageClock clk;
  // ---- a scan prefetch() handed to the worker: wait for it, the map-independent half of the step is done
  int hit = -1;
  if (!normals && workersUp_) {
    std::unique_lock<std::mutex> lock(mu_);
    for (int k = 0; k < kSets; ++k)
      if (sets_[k].state != 0 && sets_[k].srcPtr == xyz && sets_[k].srcN == n && (hit < 0 || sets_[k].ticket < sets_[hit].ticket)) hit = k;
    if (hit >= 0) {
      cv_.wait(lock, [&] { return sets_[hit].state == 2; });
      sets_[hit].state = 0;
      sets_[hit].srcPtr = nullptr;
    }
  }
  if (hit >= 0) {
    cur_ = hit;
    ScanSet& s = sets_[cur_];
    clk.lap(&stats_.prefetchWaitMs);
    stats_.prefetched = 1;
    for (int i = 0; i < 5; ++i) stats_.stageMs[i] = s.stageMs[i];   // measured on the worker, concurrent with the previous steps
    if (!s.ok) { error_ = s.error; return false; }
    return step(time, odom, true);
  }
  // ---- inline: a set no pending prefetch owns
  {
    std::unique_lock<std::mutex> lock(mu_);
    cur_ = takeSetLocked(lock);
  }
  if (b200reg_cloud64_upload(ctx_, xyz, normals, n, &sets_[cur_].raw) != B200REG_OK) { error_ = b200reg_last_error(ctx_); return false; }
  clk.lap(&stats_.stageMs[MapperStepStats::kWideCrop]);   // (the upload is booked with the first stage)
  return step(time, odom, false);
}

bool Mapper::addRangeMeasurementPointCloud2(const uint8_t* data, size_t bytes, const b200reg_pc2_layout& layout, double time, const double odom[16]) {
  stats_ = MapperStepStats();
  error_.clear();
  StageClock clk;
  {
    std::unique_lock<std::mutex> lock(mu_);
    cur_ = takeSetLocked(lock);
  }
  if (b200reg_cloud64_from_pointcloud2(ctx_, data, bytes, &layout, &sets_[cur_].raw) != B200REG_OK) { error_ = b200reg_last_error(ctx_); return false; }
  clk.lap(&stats_.stageMs[MapperStepStats::kWideCrop]);
  return step(time, odom, false);
}

bool Mapper::step(double time, const double odom[16], bool processed) {
  ScanSet& cur = sets_[cur_];
  auto processInline = [&]() {   // the scan-processing chain on this thread's context (no prefetch for this scan)
    if (processed) return true;
    for (double& v : cur.stageMs) v = 0.0;
    const bool ok = process(ctx_, cur);
    for (int i = 0; i < 5; ++i) stats_.stageMs[i] += cur.stageMs[i];
    if (!ok) error_ = cur.error;
    return ok;
  };
  // ---- first scan (Mapper.cpp:179-193)
  if (b200reg_map_size(map_) == 0) {
    std::memcpy(mapToRangeSensorPrev_, mapToRangeSensor_, sizeof(mapToRangeSensor_));
    if (!processInline()) return false;
    if (!insert()) return false;
    std::memcpy(odomPrev_, odom, sizeof(odomPrev_));   // (the reference reads its odometry buffer at the previous scan's stamp)
    haveOdomPrev_ = true;
    stats_.inserted = 1;
    stats_.mergePoints = b200reg_cloud64_size(cur.merge);
    stats_.mapPoints = b200reg_map_size(map_);
    return true;
  }
  // ---- scan-to-map initial guess: previous refined pose * odometry motion (:265-281)
  double estimate[16];
  std::memcpy(estimate, mapToRangeSensorPrev_, sizeof(estimate));
  if (haveOdomPrev_) {
    double inv[16], motion[16];
    inverseAffine(odomPrev_, inv);
    mul(inv, odom, motion);
    mul(mapToRangeSensorPrev_, motion, estimate);
  }
  // ---- scan processing and the fp64 -> fp32 conversion (:307-309)
  if (!processInline()) return false;
  stats_.readingPoints = b200reg_cloud64_size(cur.match);
  stats_.mergePoints = b200reg_cloud64_size(cur.merge);
  StageClock clk;
  float guess[16];
  for (int i = 0; i < 16; ++i) guess[i] = (float)estimate[i];
  // ---- map patch around the PREVIOUS pose (:328)
  const b200reg_crop_spec patchVol = at(params_.scanMatcherCropper, mapToRangeSensor_);
  b200reg_cloud* patch = nullptr;
  if (b200reg_map_crop(ctx_, map_, &patchVol, &patch) != B200REG_OK) { error_ = b200reg_last_error(ctx_); return false; }
  stats_.patchPoints = b200reg_cloud_size(patch);
  if (stats_.patchPoints == 0) {   // mapPatch->IsEmpty() -> return false (:330-333)
    b200reg_cloud_free(ctx_, patch);
    error_ = "map patch size is zero";
    return false;
  }
  clk.lap(&stats_.stageMs[MapperStepStats::kPatch]);
  // ---- reference refresh every referenceCloudSettingPeriod seconds of sensor time (:345-367)
  float corrected[16];
  std::memcpy(corrected, guess, sizeof(corrected));
  bool failed = false;
  if (!referenceSet_ || (time - lastReferenceInit_) >= params_.referenceCloudSettingPeriod) {
    lastReferenceInit_ = time;
    const int code = b200reg_icp_set_reference(icp_, patch);
    if (code != B200REG_OK) {   // initReference throws inside the try block -> the estimate is kept (:420-422)
      error_ = b200reg_last_error(ctx_);
      stats_.icpStatus = code;
      failed = true;
    } else {
      referenceSet_ = true;
      stats_.referenceRefreshed = 1;
    }
  }
  b200reg_cloud_free(ctx_, patch);
  clk.lap(&stats_.stageMs[MapperStepStats::kReference]);
  // ---- icp_.compute(reading, {}, guess, false) (:388-393); any runtime_error leaves the initial guess (:420-422)
  if (!failed) {
    b200reg_icp_stats st;
    float T[16];
    const int code = b200reg_icp_register(icp_, cur.reading, guess, T, &st);
    stats_.icpStatus = code;
    stats_.iterations = st.iterations;
    if (code == B200REG_OK) std::memcpy(corrected, T, sizeof(corrected));
    else error_ = b200reg_last_error(ctx_);
  }
  clk.lap(&stats_.stageMs[MapperStepStats::kRegister]);
  for (int i = 0; i < 16; ++i) mapToRangeSensor_[i] = (double)corrected[i];   // cast back to fp64 (:435)
  // ---- insertion gate (:483-489)
  double inv[16], motion[16];
  inverseAffine(lastInsertion_, inv);
  mul(inv, mapToRangeSensor_, motion);
  const double moved = std::sqrt(motion[3] * motion[3] + motion[7] * motion[7] + motion[11] * motion[11]);
  if (!(moved < params_.minMovementBetweenMappingSteps)) {
    if (!insert()) return false;
    std::memcpy(lastInsertion_, mapToRangeSensor_, sizeof(lastInsertion_));
    stats_.inserted = 1;
  }
  std::memcpy(mapToRangeSensorPrev_, mapToRangeSensor_, sizeof(mapToRangeSensor_));
  std::memcpy(odomPrev_, odom, sizeof(odomPrev_));
  stats_.mapPoints = b200reg_map_size(map_);
  return true;
}

}  // namespace b200
