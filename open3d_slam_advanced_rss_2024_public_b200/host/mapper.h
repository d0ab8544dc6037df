// host/mapper.h — the scan-to-map step of open3d_slam's Mapper, as a thin host driver over the C ABI.
//
// Mirrors the control flow of o3d_slam::Mapper::addRangeMeasurement
// (open3d_slam_rsl/open3d_slam/open3d_slam/src/Mapper.cpp:168-504) for one active submap:
//   first scan        -> preprocess + insert                                             Mapper.cpp:179-193
//   estimate          =  mapToRangeSensorPrev * (odomPrev^-1 * odomNow)                  :265-281
//   processed         =  ScanToMapIcp::processForScanMatchingAndMerging(rawScan)         ScanToMapRegistration.cpp:36-69
//                        (map-builder crop at the identity pose -> o3d_slam::voxelize -> normals if needed -> narrow crop)
//   reading           =  open3dToPointmatcher(processed.match)  fp64 -> fp32             Mapper.cpp:307-309
//   mapPatch          =  cropSubmap(activeSubmap, mapToRangeSensor_)  (PREVIOUS pose)     :328, ScanToMapRegistration.cpp:90-96
//   every `referenceCloudSettingPeriod` seconds of SENSOR time: icp.initReference(mapPatch)       :345-367
//   corrected         =  icp.compute(reading, {}, estimate, false); a runtime_error keeps the estimate  :388-422
//   insert only if the sensor moved >= minMovementBetweenMappingSteps since the last insertion     :483-489
//                        Submap::insertScan: transform, append, re-voxelise inside the cropping volume  Submap.cpp:39-96
// Out of this driver: submap switching, the odometry time buffer (the caller hands the
// odometry pose of each scan), RandomDownSample (ratio 1 only — the reference itself requires that for reproducibility,
// ScanToMapRegistration.cpp:43), the initial-map mode, motion compensation.
#pragma once
#include <condition_variable>
#include <cstdint>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200reg.h"

namespace b200 {

struct MapperParams {
  // scan processing (ScanToMapRegistration.cpp:36-69; param_*.lua: map_builder.cropper, scan_to_map_registration.scan_processing)
  b200reg_crop_spec mapBuilderCropper{};    // wide crop of the raw scan, applied at the IDENTITY pose (its pose is never set there)
  double scanVoxelSize = 0.1;               // o3d_slam::voxelize(scanProcessing_.voxelSize_), <= 0: off (helpers.cpp:109-111)
  b200reg_crop_spec scanMatcherCropper{};   // narrow crop of the processed scan (identity pose) and of the map (previous pose)
  int normalsKnn = 20;                      // CloudRegistration.cpp:62-75: EstimateNormals(Hybrid(radius, knn)) when the scan has none
  double normalsRadius = 1.0;
  // map (Submap.cpp:39-96; map_builder.*)
  double mapVoxelSize = 0.15;
  double minMovementBetweenMappingSteps = 0.0;   // Mapper.cpp:483-485
  double referenceCloudSettingPeriod = 2.0;      // seconds of sensor time, Mapper.cpp:345-349
  // space carving (Submap.cpp:56-60,130-144; map_builder.space_carving.*): off by default like Parameters.hpp:181
  bool carvingEnabled = false;
  int carveSpaceEveryNscans = 10;
  b200reg_carve_params carving{0.1, 20.0, 0.1, 0.5};
  // ICP chain (param/icp.yaml)
  b200reg_icp_config icp{};
};

struct MapperStepStats {
  int32_t iterations = 0;
  int32_t icpStatus = 0;          // b200reg status of the registration (non-zero: the estimate was kept, Mapper.cpp:420-422)
  int32_t referenceRefreshed = 0;
  int32_t inserted = 0;
  int64_t readingPoints = 0, mergePoints = 0, patchPoints = 0, mapPoints = 0, carvedPoints = 0;
  // host wall time per stage (every stage ends in a synchronising call): wide crop, voxelise, normals, narrow crop,
  // reading upload, map-patch crop, initReference, registration, insertion
  enum { kWideCrop, kVoxelize, kNormals, kNarrowCrop, kUpload, kPatch, kReference, kRegister, kInsert, kStages };
  double stageMs[kStages] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  // prefetch(): the scan-processing stages ran on the worker, concurrently with the previous step (their times are the worker's);
  // prefetchWaitMs is what this step still had to wait for them
  double prefetchWaitMs = 0;
  int32_t prefetched = 0;
};

class Mapper {
 public:
  Mapper(b200reg_ctx* ctx, const MapperParams& params);
  ~Mapper();
  Mapper(const Mapper&) = delete;
  Mapper& operator=(const Mapper&) = delete;

  // Mapper::setMapToRangeSensor(Initial) (Mapper.cpp:133-166): pose of the sensor in the map before the first scan
  void setMapToRangeSensor(const double T[16]);
  // rawScan: n x 3 fp64 points (+ optional normals) in the sensor frame; time: seconds; odomToRangeSensor: row-major 4x4 of
  // the external odometry at `time`.  Returns false where the reference returns false (empty map patch, failed initReference).
  bool addRangeMeasurement(const double* xyz, const double* normals, int64_t n, double time, const double odomToRangeSensor[16]);

  // the same step fed with the sensor_msgs/PointCloud2 bytes of the scan (SlamWrapperRos's rosToOpen3d + addRangeScan path,
  // open3d_conversions.cpp:203-244): the message is deserialised on the device, nothing else crosses PCIe
  bool addRangeMeasurementPointCloud2(const uint8_t* data, size_t bytes, const b200reg_pc2_layout& layout, double time,
                                      const double odomToRangeSensor[16]);

  // Process a coming scan AHEAD: worker threads with contexts (stream, scratch) of their own upload it and run the whole
  // scan-processing chain (wide crop -> voxelize -> normals -> narrow crop -> fp32 reading, ScanToMapRegistration.cpp:36-69 — nothing
  // in it depends on the map or the pose) while this thread registers and inserts the current scan; what SlamWrapper's scan buffer in
  // front of the mapping worker gives room for.  Call it BEFORE addRangeMeasurement of the current scan.  kWorkers scans are processed
  // concurrently and kSets - 1 may be pending; xyz must stay unchanged until the addRangeMeasurement(xyz, nullptr, n, ...) call that
  // consumes it, which waits for the worker and then only has the map-dependent half of the step left.  A measurement whose arguments
  // match no pending scan is processed inline as before.  Results are identical either way.
  bool prefetch(const double* xyz, int64_t n);

  const double* mapToRangeSensor() const { return mapToRangeSensor_; }
  b200reg_map* map() { return map_; }
  const MapperStepStats& lastStats() const { return stats_; }
  const std::string& lastError() const { return error_; }

 private:
  // ProcessedScans (ScanToMapRegistration.hpp) of ONE scan on the device: merge (wide) and match (narrow) fp64 clouds + the fp32
  // DataPoints mirror of match; kSets of them, reused scan after scan — one being consumed by the step, the others being filled ahead
  struct ScanSet {
    b200reg_cloud64 *raw = nullptr, *cropped = nullptr, *merge = nullptr, *match = nullptr;
    b200reg_cloud* reading = nullptr;   // open3dToPointmatcher(match), device to device
    const double* srcPtr = nullptr;     // host scan a pending prefetch was given
    int64_t srcN = -1;
    int state = 0;                      // 0 free / consumed, 1 queued or being processed by the worker, 2 ready
    uint64_t ticket = 0;                // order of the prefetch calls (the older pending scan is replaced first)
    bool ok = false;
    std::string error;
    double stageMs[5] = {0, 0, 0, 0, 0};   // wide crop (+ upload), voxelize, normals, narrow crop, fp32 conversion
  };
  bool process(b200reg_ctx* ctx, ScanSet& s);   // s.raw -> s.merge / s.match / s.reading
  bool step(double time, const double odomToRangeSensor[16], bool processed);
  bool insert();
  void workerLoop(int w);
  void freeSet(ScanSet& s);
  int takeSetLocked(std::unique_lock<std::mutex>& lock);   // a free set, else the oldest pending one once its worker is done with it
  static constexpr int kSets = 3, kWorkers = 2;

  b200reg_ctx* ctx_;
  MapperParams params_;
  b200reg_map* map_ = nullptr;
  b200reg_icp* icp_ = nullptr;
  ScanSet sets_[kSets];
  int cur_ = 0;                        // the set the current step works on
  // the workers of prefetch(): created on first use
  b200reg_ctx* workerCtx_[kWorkers] = {nullptr, nullptr};
  std::thread worker_[kWorkers];
  bool workersUp_ = false;
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<int> jobs_;               // sets waiting for a worker, in prefetch order
  bool stop_ = false;
  uint64_t tickets_ = 0;
  double mapToRangeSensor_[16], mapToRangeSensorPrev_[16], lastInsertion_[16], odomPrev_[16];
  double cropperPose_[16];             // mapBuilderCropper_'s pose: set AFTER each insertion (Submap.cpp:86), so carving sees the previous one
  int64_t nScansInsertedMap_ = 0;
  bool haveOdomPrev_ = false, referenceSet_ = false;
  double lastReferenceInit_ = 0.0;
  MapperStepStats stats_;
  std::string error_;
};

}  // namespace b200
