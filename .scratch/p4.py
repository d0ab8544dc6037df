p='open3d_slam_advanced_rss_2024_public_b200/host/mapper.cpp'
s=open(p).read()
i0=s.index("// ScanToMapIcp::processForScanMatchingAndMerging (ScanToMapRegistration.cpp:36-69)")
i1=s.index("bool Mapper::addRangeMeasurement")
new=r'''// ScanToMapIcp::processForScanMatchingAndMerging (ScanToMapRegistration.cpp:36-69), every step on the device: the raw scan is
// the only host -> device copy of a mapping step
bool Mapper::process(const double* xyz, const double* normals, int64_t n) {
  StageClock clk;
  auto fail = [&]() { error_ = b200reg_last_error(ctx_); return false; };
  if (b200reg_cloud64_upload(ctx_, xyz, normals, n, &raw_) != B200REG_OK) return fail();
  // wide crop: mapBuilderCropper_->crop(in) — that cropper instance never gets a pose, i.e. it sits at the identity (:30-37)
  const b200reg_crop_spec wide = at(params_.mapBuilderCropper, nullptr);
  if (b200reg_cloud64_crop(ctx_, &wide, raw_, &cropped_) != B200REG_OK) return fail();
  clk.lap(&stats_.stageMs[MapperStepStats::kWideCrop]);
  // o3d_slam::voxelize (helpers.cpp:108-115): no-op for voxel <= 0
  if (params_.scanVoxelSize > 0.0 && b200reg_cloud64_size(cropped_) > 0) {
    if (b200reg_cloud64_voxel_down_sample(ctx_, params_.scanVoxelSize, cropped_, &merge_) != B200REG_OK) return fail();
  } else {
    std::swap(cropped_, merge_);
    if (!merge_) { error_ = "ScanToMapIcp::wideCropped cropped size is zero"; return false; }
  }
  if (b200reg_cloud64_size(merge_) == 0) { error_ = "ScanToMapIcp::wideCropped cropped size is zero"; return false; }   // assert_gt (:66)
  clk.lap(&stats_.stageMs[MapperStepStats::kVoxelize]);
  // estimateNormalsOrCovariancesIfNeeded (CloudRegistration.cpp:62-75): only when the cloud has no normals;
  // EstimateNormals(Hybrid(radius, knn)) + NormalizeNormals + OrientNormalsTowardsCameraLocation(origin)
  if (!b200reg_cloud64_has_normals(merge_)) {
    const double origin[3] = {0.0, 0.0, 0.0};
    int32_t degenerate = 0;
    if (b200reg_cloud64_estimate_normals(ctx_, merge_, params_.normalsKnn, params_.normalsRadius, origin, &degenerate) != B200REG_OK) return fail();
  }
  clk.lap(&stats_.stageMs[MapperStepStats::kNormals]);
  // narrow crop at the identity pose (:62-63)
  const b200reg_crop_spec narrow = at(params_.scanMatcherCropper, nullptr);
  if (b200reg_cloud64_crop(ctx_, &narrow, merge_, &match_) != B200REG_OK) return fail();
  if (b200reg_cloud64_size(match_) == 0) { error_ = "ScanToMapIcp::narrow cropped size is zero"; return false; }   // assert_gt (:65)
  clk.lap(&stats_.stageMs[MapperStepStats::kNarrowCrop]);
  return true;
}

// Submap::insertScan (Submap.cpp:39-96) without carving: transform, append, re-voxelise inside the map-builder volume at the pose
bool Mapper::insert() {
  StageClock clk;
  const b200reg_crop_spec vol = at(params_.mapBuilderCropper, mapToRangeSensor_);
  if (b200reg_map_insert_cloud64(ctx_, map_, merge_, mapToRangeSensor_, &vol) != B200REG_OK) {
    error_ = b200reg_last_error(ctx_);
    return false;
  }
  clk.lap(&stats_.stageMs[MapperStepStats::kInsert]);
  return true;
}

'''
s=s[:i0]+new+s[i1:]
s=s.replace("""  Processed processed;
""","")
s=s.replace("""    if (!process(xyz, normals, n, &processed)) return false;
    if (!insert(processed)) return false;""","""    if (!process(xyz, normals, n)) return false;
    if (!insert()) return false;""")
s=s.replace("""    stats_.mergePoints = (int64_t)(processed.mergeXyz.size() / 3);
    stats_.mapPoints = b200reg_map_size(map_);
    return true;
  }""","""    stats_.mergePoints = b200reg_cloud64_size(merge_);
    stats_.mapPoints = b200reg_map_size(map_);
    return true;
  }""")
s=s.replace("""  if (!process(xyz, normals, n, &processed)) return false;
  const int64_t nRead = (int64_t)(processed.matchXyz.size() / 3);
  stats_.readingPoints = nRead;
  stats_.mergePoints = (int64_t)(processed.mergeXyz.size() / 3);
  StageClock clk;
  if (reading_) { b200reg_cloud_free(ctx_, reading_); reading_ = nullptr; }
  if (b200reg_cloud_upload_f64(ctx_, processed.matchXyz.data(), processed.matchNrm.data(), nRead, &reading_) != B200REG_OK) {
    error_ = b200reg_last_error(ctx_);
    return false;
  }""","""  if (!process(xyz, normals, n)) return false;
  stats_.readingPoints = b200reg_cloud64_size(match_);
  stats_.mergePoints = b200reg_cloud64_size(merge_);
  StageClock clk;
  if (b200reg_cloud64_to_f32(ctx_, match_, &reading_) != B200REG_OK) {   // open3dToPointmatcher, device to device
    error_ = b200reg_last_error(ctx_);
    return false;
  }""")
s=s.replace("    if (!insert(processed)) return false;\n    std::memcpy(lastInsertion_","    if (!insert()) return false;\n    std::memcpy(lastInsertion_")
s=s.replace("""Mapper::~Mapper() {
  if (reading_) b200reg_cloud_free(ctx_, reading_);""","""Mapper::~Mapper() {
  if (reading_) b200reg_cloud_free(ctx_, reading_);
  b200reg_cloud64_free(ctx_, raw_);
  b200reg_cloud64_free(ctx_, cropped_);
  b200reg_cloud64_free(ctx_, merge_);
  b200reg_cloud64_free(ctx_, match_);""")
s=s.replace("#include <cstring>\n","#include <cstring>\n#include <utility>\n")
open(p,'w').write(s)
p='open3d_slam_advanced_rss_2024_public_b200/host/mapper.h'
s=open(p).read()
s=s.replace("""  struct Processed {   // ProcessedScans (ScanToMapRegistration.hpp): match_ (narrow) and merge_ (wide), fp64 host clouds
    std::vector<double> mergeXyz, mergeNrm, matchXyz, matchNrm;
  };
  bool process(const double* xyz, const double* normals, int64_t n, Processed* out);
  bool insert(const Processed& p);
""","""  // ProcessedScans (ScanToMapRegistration.hpp): merge_ (wide) and match_ (narrow), device-resident fp64 clouds reused scan after scan
  bool process(const double* xyz, const double* normals, int64_t n);
  bool insert();
""")
s=s.replace("""  b200reg_cloud* reading_ = nullptr;   // reused device cloud of the reading (streaming upload)
""","""  b200reg_cloud* reading_ = nullptr;   // fp32 DataPoints mirror of match_ (open3dToPointmatcher on the device)
  b200reg_cloud64 *raw_ = nullptr, *cropped_ = nullptr, *merge_ = nullptr, *match_ = nullptr;
""")
open(p,'w').write(s)
print('ok')
