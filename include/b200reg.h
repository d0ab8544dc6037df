/*
 * include/b200reg.h — C ABI of libb200reg.so: the B200 (sm_100a) scan-to-map registration hot path.
 *
 * Drop-in boundary for the path the reference implements in libpointmatcher's registrable modules and open3d_slam's
 * Mapper.  Plain pointers and sizes only; no C++ types, no exceptions, no torch types.  Each entry
 * point cites the reference interface it replaces (paths relative to /root/reference/).
 *
 * Conventions
 *   - every function returns int32_t status: B200REG_OK (0) or a negative B200REG_E_* code; b200reg_last_error(ctx)
 *     returns the message the C++ shim re-throws as the matching libpointmatcher exception type.
 *   - one ctx per (thread, device); calls on one ctx are ordered on its CUDA stream; a ctx is NOT thread-safe,
 *     distinct ctxs are independent (libpointmatcher modules are stateful and non re-entrant too).
 *   - host cloud layout == Eigen column-major DataPoints (PointMatcher.h:229-431): features = `feat_stride` floats per
 *     point with x y z first (stride 4: x y z pad), normals = 3 floats per point.  Copy semantics; caller owns host memory.
 *   - matches (PointMatcher.h:444-464): ids[k*i + j], d2[k*i + j] = j-th neighbour of reading point i, SQUARED distances,
 *     invalid = (-1, +inf).
 *   - 4x4 transforms are ROW-MAJOR float[16] / double[16] at this boundary.
 *   - there is NO CPU fallback: without a CUDA device every call fails with B200REG_E_CUDA.
 *   - device memory comes from the device's DEFAULT stream-ordered pool (cudaMallocAsync); b200reg_ctx_create raises that pool's
 *     release threshold to "never" so that freed blocks stay cached — a process-wide setting of the device the host application
 *     shares (cudaMemPoolTrimTo on cudaDeviceGetDefaultMemPool gives the memory back).
 *   - env B200REG_DEBUG_SYNC=1 synchronises after every checked kernel launch, so a faulting kernel is reported at its own call.
 */
#ifndef B200REG_H
#define B200REG_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define B200REG_ABI_VERSION 2

enum {
  B200REG_OK = 0,
  B200REG_E_INVALID_PARAM = -1,   /* PointMatcherSupport::InvalidParameter (Parametrizable.h)            */
  B200REG_E_INVALID_FIELD = -2,   /* DataPoints::InvalidField (PointMatcher.h:304-307)                    */
  B200REG_E_NO_POINTS = -3,       /* runtime_error "The reading point cloud is empty" (ICP.cpp:357-359)   */
  B200REG_E_NOT_CONVERGED = -4,   /* PointMatcher::ConvergenceError (PointMatcher.h:100-103)              */
  B200REG_E_TRANSFORM = -5,       /* TransformationError (TransformationsImpl.cpp:74)                     */
  B200REG_E_CUDA = -6,
  B200REG_E_OOM = -7
};

typedef struct b200reg_ctx b200reg_ctx;
typedef struct b200reg_cloud b200reg_cloud;   /* device-resident DataPoints mirror: float4 xyz(+pad), float4 normal */
typedef struct b200reg_index b200reg_index;   /* radix-sorted voxel-grid index over a (centred) reference cloud    */
typedef struct b200reg_map b200reg_map;       /* resident fp64 voxel map (open3d_slam Submap::mapCloud_)           */
typedef struct b200reg_cloud64 b200reg_cloud64; /* device-resident open3d::geometry::PointCloud mirror: fp64 xyz (+ normals) */

/* ---- context ------------------------------------------------------------------------------------------------ */
int32_t b200reg_abi_version(void);
int32_t b200reg_ctx_create(int32_t device, b200reg_ctx** out);
void    b200reg_ctx_destroy(b200reg_ctx* ctx);
const char* b200reg_last_error(const b200reg_ctx* ctx);
int32_t b200reg_ctx_sync(b200reg_ctx* ctx);
int32_t b200reg_ctx_device(const b200reg_ctx* ctx);   /* the CUDA device ordinal the context was created on */
/* number of kernels this ctx has launched since creation (bench.py "gpu_launches") */
uint64_t b200reg_ctx_launch_count(const b200reg_ctx* ctx);

/* measurement helpers (bench.py): CUDA-event timer on the ctx stream, and an L2 flush (writes a buffer larger than L2) */
int32_t b200reg_ctx_timer_start(b200reg_ctx* ctx);
int32_t b200reg_ctx_timer_stop(b200reg_ctx* ctx, float* elapsed_ms);   /* synchronises on the stop event */
int32_t b200reg_ctx_flush_l2(b200reg_ctx* ctx);
/* kernel-only timing of the match kernel (the roofline's dominant kernel): while profiling is on, every k-NN launch on this
 * ctx — standalone or inside b200reg_icp_register — is bracketed by its own CUDA event pair on the ctx stream (up to 256
 * launches; the graph-driven registration loop is replaced by the host-driven one so the brackets exist).
 * profile_read synchronises, writes the per-launch durations in launch order, and clears the list. */
int32_t b200reg_ctx_set_profiling(b200reg_ctx* ctx, int32_t on);
int32_t b200reg_ctx_profile_read(b200reg_ctx* ctx, float* ms_out, int32_t cap, int32_t* count);
/* device-side trace of the graph-resident registration loop (a profiler does not see inside a conditional graph node): while
 * on, the first thread of every kernel of the loop stamps %globaltimer.  trace_read returns the stamps of the LAST
 * registration: (kernel tag << 56) | nanoseconds; tags 1 match, 2-4 select passes, 5 reduce, 6 tail start, 7 tail end. */
int32_t b200reg_ctx_set_trace(b200reg_ctx* ctx, int32_t on);
int32_t b200reg_ctx_trace_read(b200reg_ctx* ctx, uint64_t* out, int32_t cap, int32_t* count);

/* ---- clouds: replaces DataPoints construction/copies (PointMatcher.h:229-431) and open3dToPointmatcher
 *      (open3d_slam_rsl/open3d_utils/open3d_conversions/src/open3d_conversions.cpp:21-118) -------------------------- */
int32_t b200reg_cloud_upload_f32(b200reg_ctx* ctx, const float* feat, int32_t feat_stride, const float* normals, int64_t n,
                                 b200reg_cloud** out);
/* streaming re-fill of an existing device cloud (no allocation once its capacity suffices).  async != 0: enqueue only; the host
 * arrays must stay valid and unmodified until the next synchronising call on this ctx (pinned memory: upload + registration = one sync) */
int32_t b200reg_cloud_update_f32(b200reg_ctx* ctx, b200reg_cloud* cloud, const float* feat, int32_t feat_stride, const float* normals,
                                 int64_t n, int32_t async);
int32_t b200reg_cloud_upload_f64(b200reg_ctx* ctx, const double* xyz, const double* normals, int64_t n, b200reg_cloud** out);
int32_t b200reg_cloud_download_f32(b200reg_ctx* ctx, const b200reg_cloud* cloud, float* feat, int32_t feat_stride, float* normals);
int64_t b200reg_cloud_size(const b200reg_cloud* cloud);
int32_t b200reg_cloud_has_normals(const b200reg_cloud* cloud);
void    b200reg_cloud_free(b200reg_ctx* ctx, b200reg_cloud* cloud);

/* ---- RigidTransformation::inPlaceCompute (libpointmatcher/pointmatcher/TransformationsImpl.cpp:61-114) -------------- */
int32_t b200reg_cloud_transform(b200reg_ctx* ctx, b200reg_cloud* inout, const float T[16]);

/* ---- KDTreeMatcher::init (MatchersImpl.cpp:109-114) -> index build; KDTreeMatcher::findClosests (:117-132) -> knn ---- */
/* cell <= 0 : automatic cell size.  The index copies what it needs; `ref` may be freed afterwards.  The directory holds at most
 * max(8 x points, 65536) cells (<= 2^30): a smaller requested cell is grown by factors of 1.26 until the grid fits (the search stays
 * exact for any cell size).  Non-finite coordinates in `ref` -> B200REG_E_INVALID_PARAM. */
int32_t b200reg_index_build(b200reg_ctx* ctx, const b200reg_cloud* ref, float cell, b200reg_index** out);
void    b200reg_index_free(b200reg_ctx* ctx, b200reg_index* index);
int64_t b200reg_index_size(const b200reg_index* index);
float   b200reg_index_cell(const b200reg_index* index);
/* T (nullable): applied to every query on the fly (fp32, fixed op order) before the search — the fused
 * `transformations.apply(stepReading, T_iter)` + `matcher->findClosests(stepReading)` of ICP.cpp:409-413.
 * ids_out / d2_out : HOST buffers of k*n entries (nullable).  visit_count (nullable) = points distance-tested
 * (Matcher::visitCounter, Matcher.cpp:46-72).  epsilon is accepted for YAML compatibility and treated as 0. */
int32_t b200reg_knn(b200reg_ctx* ctx, const b200reg_index* index, const b200reg_cloud* query, const float* T, int32_t k,
                    float max_dist, float epsilon, int32_t allow_self, int32_t* ids_out, float* d2_out, uint64_t* visit_count);

/* ---- one ICP iteration's device work (ICP.cpp:409-433 minus the 6x6 solve):
 *      transform + match (a4,a6), TrimmedDistOutlierFilter (OutlierFiltersImpl.cpp:140-147 + Matches.cpp:61-87),
 *      SurfaceNormalOutlierFilter (OutlierFiltersImpl.cpp:226-281), ErrorElements statistics (ErrorMinimizer.cpp:59-140)
 *      and the point-to-plane normal equations A = G G^T, b = -G h^T (PointToPlane.cpp:107-156) ---------------------- */
typedef struct b200reg_step_stats {
  int64_t kept;                       /* pairs with non-zero weight and finite distance               */
  int64_t n_finite;                   /* finite matches (input of the quantile)                       */
  double  weight_sum;                 /* sum of weights                                              */
  float   tau;                        /* trimmed-distance limit (bit-exact nth_element value); inf if filter off */
  float   point_used_ratio;           /* ErrorMinimizer.cpp:139                                      */
  float   weighted_point_used_ratio;  /* ErrorMinimizer.cpp:140                                      */
  uint64_t visit_count;
  double  mean_reading[3];            /* mean of kept (transformed) reading points  (PointToPlane.cpp:263) */
  double  mean_reference[3];          /* mean of matched reference points           (PointToPlane.cpp:264) */
} b200reg_step_stats;

/* A_out: 6x6 row-major (unknown order alpha beta gamma tx ty tz), formed on the mean-centred pairs exactly as
 * PointToPlane.cpp:263-268 does; b_out: 6.  trim_ratio <= 0 / max_angle < 0 switch the respective filter off
 * (both off = OutlierFilters::compute with an empty chain, OutlierFilter.cpp:70-85).  force2D != 0 drops the z components
 * of points and normals before forming G and h (PointToPlane.cpp:247-254). */
int32_t b200reg_icp_step(b200reg_ctx* ctx, const b200reg_index* index, const b200reg_cloud* reading, const float* T_iter,
                         int32_t k, float max_dist, float trim_ratio, float max_angle, int32_t force2D, double A_out[36],
                         double b_out[6], b200reg_step_stats* stats);

/* ---- the same kernels behind the individual modules' Eigen-shaped interfaces (HOST arrays in / out): what the C++ shim
 *      classes call when a chain mixes CPU and CUDA modules --------------------------------------------------------- */
/* TrimmedDistOutlierFilter::compute (OutlierFiltersImpl.cpp:140-147): tau = quantile of the finite d2, w = (d2 <= tau) */
int32_t b200reg_trimmed_dist(b200reg_ctx* ctx, const float* d2, int64_t count, float ratio, float* tau_out, float* w_out);
/* MaxDistOutlierFilter / MinDistOutlierFilter / MedianDistOutlierFilter::compute (OutlierFiltersImpl.cpp:64-125) on the squared
 * distances of a Matches object: w = (d2 <= maxDist^2), (d2 >= minDist^2), (d2 <= factor * median(d2)) with the median taken by
 * the same exact radix select as the trimmed filter (Matches::getDistsQuantile(0.5)).  param = maxDist | minDist | factor;
 * limit_out (nullable) = the threshold that was applied. */
enum { B200REG_OUTLIER_MAX_DIST = 0, B200REG_OUTLIER_MIN_DIST = 1, B200REG_OUTLIER_MEDIAN_DIST = 2 };
int32_t b200reg_dist_outlier(b200reg_ctx* ctx, int32_t kind, const float* d2, int64_t count, float param, float* limit_out, float* w_out);
/* VarTrimmedDistOutlierFilter::compute (OutlierFiltersImpl.cpp:166-220): the inlier ratio that minimises FRMS(f) =
 * cumsum(sorted d2)[f] / f / (f / N)^(2 lambda) over [minRatio, maxRatio) (sorted = finite d2 > 0; the fp32 running sum is
 * taken in the reference's order), then limit = getDistsQuantile(ratio), w = (d2 <= limit).  ratio_out / tau_out / w_out nullable. */
int32_t b200reg_var_trimmed_dist(b200reg_ctx* ctx, const float* d2, int64_t count, float min_ratio, float max_ratio, float lambda,
                                 float* ratio_out, float* tau_out, float* w_out);
/* SurfaceNormalOutlierFilter::compute (OutlierFiltersImpl.cpp:236-281); normals 3 floats per point, nullable = filter off */
int32_t b200reg_surface_normal_outlier(b200reg_ctx* ctx, const float* normals_reading, const float* normals_reference,
                                       int64_t n_reference, const int32_t* ids, int32_t k, int64_t n, float max_angle, float* w_out);
/* ErrorMinimizer::compute(reading, reference, weights, matches) (ErrorMinimizer.cpp:218-232 + PointToPlane.cpp:107-156):
 * explicit matches and OutlierWeights (k*n each, HOST) over two device clouds; reference in its ORIGINAL order. */
int32_t b200reg_point_to_plane(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference, const int32_t* ids,
                               const float* d2, const float* w, int32_t k, int32_t force2D, double A_out[36], double b_out[6],
                               b200reg_step_stats* stats);

/* PointToPointErrorMinimizer::compute (ErrorMinimizers/PointToPoint.cpp:55-94) on explicit matches and OutlierWeights (k*n each,
 * HOST): weighted means, 3x3 cross-covariance of the centred pairs (fp64 on the device), SVD on the host, reflection fix,
 * T_out = [U V^T | meanReference - R meanReading] rounded to fp32.  stats nullable. */
int32_t b200reg_point_to_point(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference, const int32_t* ids,
                               const float* d2, const float* w, int32_t k, float T_out[16], b200reg_step_stats* stats);
/* PointToPlaneWithCovErrorMinimizer::estimateCovariance (ErrorMinimizers/PointToPlaneWithCov.cpp:70-160): covariance of the step
 * transform T (the one b200reg_solve_point_to_plane returned for the same pairs) from the pairs centred on mean_reading /
 * mean_reference (b200reg_step_stats of b200reg_point_to_plane) — the state PointToPlane.cpp:263-268 leaves the ErrorElements in.
 * cov_out: 6x6 row-major, sensor_std_dev^2 * H^-1 (D D^T) H^-1. */
int32_t b200reg_point_to_plane_covariance(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference, const int32_t* ids,
                                          const float* d2, const float* w, int32_t k, const double mean_reading[3],
                                          const double mean_reference[3], const float T[16], float sensor_std_dev, double cov_out[36]);

/* PointToPlaneErrorMinimizer::getResidualError / computeResidualError (ErrorMinimizers/PointToPlane.cpp:372-422): over the ErrorElements
 * pairs (finite distance, non-zero weight) the sum of w ((reading - reference) . n_reference)^2, accumulated in fp64; force2D drops z.
 * (getOverlap, :426-521: the device clouds carry no simpleSensorNoise / densities descriptors, for which the reference itself returns
 * getWeightedPointUsedRatio() (:477-481) — b200reg_step_stats.weighted_point_used_ratio / b200reg_icp_stats.overlap_ratio.) */
int32_t b200reg_point_to_plane_residual(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference, const int32_t* ids,
                                        const float* d2, const float* w, int32_t k, int32_t force2D, double* residual_out, int64_t* kept_out);

/* The same computation for callers that may not allocate (CUDA-graph capture set-ups, tight serving loops, harnesses that forbid
 * cudaMalloc).  This entry point has NO counterpart in the reference.
 * Arguments, result and error codes are those of b200reg_point_to_plane_residual (same kernels, same grid: the same bits), but the
 * device staging of ids / d2 / w and the partial sums live in a static device array of the library — 12 bytes per pair for at most
 * B200REG_NOALLOC_MAX_PAIRS pairs (reading size x k) plus 8 KiB, held by every GPU that loads the library until the process exits —
 * and the call performs no cudaMalloc / cudaFree / pool or pinned allocation.  More pairs than the maximum: B200REG_E_INVALID_PARAM
 * with a message naming the limit.  Every call on a device shares the one array (it is overwritten, and its reduction slots are
 * cleared, at the start of each call): calls that overlap on the GPU — other contexts or threads on the same device — race; calls
 * on one context are ordered by its stream. */
#define B200REG_NOALLOC_MAX_PAIRS (1 << 20)
int32_t b200reg_point_to_plane_residual_noalloc(b200reg_ctx* ctx, const b200reg_cloud* reading, const b200reg_cloud* reference,
                                                const int32_t* ids, const float* d2, const float* w, int32_t k, int32_t force2D,
                                                double* residual_out, int64_t* kept_out);

/* PointToPlaneErrorMinimizer::compute tail (PointToPlane.cpp:185-368): solve (LLT, rank-reduced minimum norm when
 * singular), Rodrigues, un-centre.  Host arithmetic in fp64; T_out = the 4x4 step transform rounded to fp32. */
int32_t b200reg_solve_point_to_plane(const double A[36], const double b[6], const double mean_reading[3],
                                     const double mean_reference[3], int32_t force2D, int32_t force4DOF, float T_out[16],
                                     double x_out[6]);

/* ---- whole registration: ICP::initReference + ICP::computeWithTransformedReference (ICP.cpp:258-468) ---------------- */
typedef struct b200reg_icp_config {
  int32_t knn;            /* KDTreeMatcher knn (MatchersImpl.h:97)                                   */
  float   max_dist;       /* KDTreeMatcher maxDist (MatchersImpl.h:100), inf allowed                 */
  float   epsilon;        /* accepted, treated as 0                                                  */
  float   trim_ratio;     /* TrimmedDistOutlierFilter ratio (OutlierFiltersImpl.h:137); <= 0 absent   */
  float   max_angle;      /* SurfaceNormalOutlierFilter maxAngle; < 0 absent                          */
  int32_t force2D;        /* PointToPlane.h:74                                                       */
  int32_t force4DOF;      /* PointToPlane.h:75                                                       */
  int32_t max_iter;       /* CounterTransformationChecker maxIterationCount; < 0 absent (0 stops after the first
                             iteration, exactly like 1: counter >= limit, TransformationCheckersImpl.cpp:70)       */
  float   min_diff_rot;   /* DifferentialTransformationChecker (TransformationCheckersImpl.h:88-93)  */
  float   min_diff_trans;
  int32_t smooth_length;  /* < 0 : checker absent                                                    */
  float   bound_rot;      /* BoundTransformationChecker; < 0 absent                                  */
  float   bound_trans;
  float   cell;           /* index cell size for b200reg_icp_set_reference; <= 0 automatic           */
  int32_t checker_order[3]; /* order the chain lists its checkers in (B200REG_CHECKER_*, 0 ends the list): the reference
                             evaluates them in that order and Counter's MaxNumIterationsReached throw skips the ones
                             behind it (ICP.cpp:437-446).  All 0 = Differential, Bound, Counter.  Whatever the
                             configuration, the loop gives up with B200REG_E_NOT_CONVERGED after 10000 iterations.    */
} b200reg_icp_config;
enum { B200REG_CHECKER_COUNTER = 1, B200REG_CHECKER_DIFFERENTIAL = 2, B200REG_CHECKER_BOUND = 3 };

typedef struct b200reg_icp_stats {   /* the reference's Inspector stat names (ICP.cpp:282-284,391-393,452-456) */
  int32_t iterations;                 /* IterationsCount                 */
  int32_t max_iterations_reached;     /* ICP.cpp:444                     */
  int64_t reference_point_count;      /* ReferencePointCount             */
  int64_t reading_point_count;        /* ReadingPointCount               */
  uint64_t points_touched;            /* PointCountTouched               */
  float   overlap_ratio;              /* OverlapRatio = weightedPointUsedRatio */
  float   point_used_ratio;
  float   last_tau;
  int64_t kept_last;
  double  reference_preprocessing_s;  /* ReferencePreprocessingDuration  */
  double  reading_preprocessing_s;    /* ReadingPreprocessingDuration    */
  double  convergence_s;              /* ConvergenceDuration             */
  double  device_ms;                  /* CUDA-event time of the iterate loop */
} b200reg_icp_stats;

typedef struct b200reg_icp b200reg_icp;   /* PointMatcher<float>::ICP with the CUDA chain inside */
int32_t b200reg_icp_create(b200reg_ctx* ctx, const b200reg_icp_config* cfg, b200reg_icp** out);
void    b200reg_icp_destroy(b200reg_icp* icp);
/* ICP::initReference (ICP.cpp:293-328): mean, centring, index build.  Keeps its own copy of what it needs.  The work is only
 * ENQUEUED on the ctx stream (mean, bounding box and grid are derived on the device): the call does not wait for the GPU, and a
 * reference with non-finite coordinates is reported (B200REG_E_INVALID_PARAM) by the next call that needs the result —
 * b200reg_icp_register, b200reg_icp_reference_mean or b200reg_index_cell(b200reg_icp_index(icp)). */
int32_t b200reg_icp_set_reference(b200reg_icp* icp, const b200reg_cloud* reference);
/* ICP::compute(reading, <empty>, T_init, initializeMatcher=false) (ICP.cpp:258-290 + 332-468), as Mapper.cpp:391-393 calls it */
int32_t b200reg_icp_register(b200reg_icp* icp, const b200reg_cloud* reading, const float T_init[16], float T_out[16],
                             b200reg_icp_stats* stats);
/* ---- batched registration: n independent pairs, pair i = initReference(refs[i]) + compute(readings[i], T_init + 16 i) -> T_out + 16 i
 *      (the loop-closure refinement loop, open3d_slam/src/PlaceRecognition.cpp:71-174).  A batch object owns `lanes`
 *      independent lanes on the ctx's device (stream, scratch, ICP object with its instantiated loop graph, reusable clouds); pair i
 *      runs on lane i mod lanes and up to `lanes` pairs are in flight at once — uploads, index build, reading preparation, the
 *      device-resident loop and the read-back of one pair are enqueued without a host wait, and the kernels of different lanes
 *      interleave on the device.  Every pair's result is bit-identical to b200reg_icp_set_reference + b200reg_icp_register.
 *      lanes <= 0: default (8, or env B200REG_BATCH_LANES).  The configuration must have knn == 1.
 *      stats / status: arrays of n, nullable.  With `status` the call returns OK and reports per-pair codes (a failed pair keeps
 *      T_out = T_init, what Mapper.cpp:420-428 falls back to); without it the first failure code is returned.
 *      One batch per (thread, device); sharding pairs over GPUs and gathering the results: b200reg_multi_* below. */
typedef struct b200reg_batch b200reg_batch;
int32_t b200reg_batch_create(b200reg_ctx* ctx, const b200reg_icp_config* cfg, int32_t lanes, b200reg_batch** out);
void    b200reg_batch_destroy(b200reg_batch* batch);
int32_t b200reg_batch_lanes(const b200reg_batch* batch);
/* ICP::initReference ONCE for the whole batch: every pair whose reference is not given (refs == NULL below) registers its reading
 * against this index — many scans, or many initial guesses of one scan, against one map (re-localisation, multi-sequence mapping). */
int32_t b200reg_batch_set_reference(b200reg_batch* batch, const b200reg_cloud* reference);
/* clouds already on the device (created on `ctx`); refs == NULL: the reference of b200reg_batch_set_reference for every pair */
int32_t b200reg_batch_register(b200reg_batch* batch, int32_t n, const b200reg_cloud* const* refs, const b200reg_cloud* const* readings,
                               const float* T_init, float* T_out, b200reg_icp_stats* stats, int32_t* status);
/* HOST clouds (layout of b200reg_cloud_upload_f32; normals nullable for readings, required for references; refs == NULL as above):
 * the uploads are part of the pipeline — every lane has a copy stream and two sets of device clouds, and the clouds of a lane's NEXT pair
 * are uploaded while its current pair iterates.  The buffers must stay valid until the call returns; pinned memory makes the copies truly
 * asynchronous (pageable memory is staged by the runtime and blocks the calling thread for the staging copy only). */
typedef struct b200reg_host_cloud {
  const float* feat;
  int32_t feat_stride;
  const float* normals;
  int64_t n;
} b200reg_host_cloud;
int32_t b200reg_batch_register_host(b200reg_batch* batch, int32_t n, const b200reg_host_cloud* refs, const b200reg_host_cloud* readings,
                                    const float* T_init, float* T_out, b200reg_icp_stats* stats, int32_t* status);
/* one-shot form of b200reg_batch_register: a batch object with the default number of lanes for the duration of the call */
int32_t b200reg_icp_register_batch(b200reg_ctx* ctx, const b200reg_icp_config* cfg, int32_t n, const b200reg_cloud* const* refs,
                                   const b200reg_cloud* const* readings, const float* T_init, float* T_out, b200reg_icp_stats* stats,
                                   int32_t* status);
/* ---- the batch sharded over the GPUs of one box from ONE process: pair i -> devices[i mod G]; every device runs its pairs
 *      through its own b200reg_batch, driven by its own host thread; then ONE ncclAllGather of 96-byte records (4x4, status,
 *      iterations, kept, tau, owner) over NVLink leaves every device with every pair's result, and T_out / status are read from
 *      the gathered block.  NCCL is resolved at run time and used for this gather only: the path in env B200REG_NCCL_LIB, else an NCCL
 *      the process already holds, else the system's libnccl.so.2.  (A process keeps ONE libnccl.so.2: an application that loads a newer
 *      NCCL under that soname later — `import torch` with its bundled copy — must load it before the first b200reg_multi_create.)
 *      devices == NULL / n_devices == 0: every visible GPU.  On failure *out is still returned so that b200reg_multi_last_error can be
 *      read; destroy it.  gather_ms (nullable): CUDA-event time of the all-gather on device 0. */
typedef struct b200reg_multi b200reg_multi;
int32_t b200reg_multi_create(const int32_t* devices, int32_t n_devices, const b200reg_icp_config* cfg, int32_t lanes, b200reg_multi** out);
void    b200reg_multi_destroy(b200reg_multi* multi);
int32_t b200reg_multi_devices(const b200reg_multi* multi);
int32_t b200reg_multi_nccl_version(const b200reg_multi* multi);
const char* b200reg_multi_last_error(const b200reg_multi* multi);
uint64_t b200reg_multi_launch_count(const b200reg_multi* multi);
int32_t b200reg_multi_register_host(b200reg_multi* multi, int32_t n, const b200reg_host_cloud* refs, const b200reg_host_cloud* readings,
                                    const float* T_init, float* T_out, b200reg_icp_stats* stats, int32_t* status, float* gather_ms);
const b200reg_index* b200reg_icp_index(const b200reg_icp* icp);
int32_t b200reg_icp_reference_mean(const b200reg_icp* icp, float mean_out[3]);

/* ---- DataPointsFilters -------------------------------------------------------------------------------------------- */
/* VoxelGridDataPointsFilter::inPlaceFilter (DataPointsFilters/VoxelGrid.cpp:71-344); origin_mode 0 = libpointmatcher
 * bbox convention.  use_centroid 0 = cell centre (corrected).  Normals are averaged like any descriptor
 * when average_descriptors != 0, else the first point's are kept. */
int32_t b200reg_voxelgrid(b200reg_ctx* ctx, b200reg_cloud* inout, float vx, float vy, float vz, int32_t use_centroid,
                          int32_t average_descriptors);
/* The compacting filters: MaxDistDataPointsFilter (MaxDist.cpp:62-105: keep norm < |maxDist|, or feature(dim) < maxDist),
 * MinDistDataPointsFilter (MinDist.cpp:62-101, the same with >), BoundingBoxDataPointsFilter (BoundingBox.cpp:66-95, strict
 * bounds, removeInside) and RandomSamplingDataPointsFilter (RandomSampling.cpp:60-100: draw i of std::minstd_rand(seed) against
 * prob, at most floor(n * prob) + 1 points; the reference seeds from std::random_device, here the seed is the caller's).
 * Order of the kept points is preserved; normals travel with their points. */
enum { B200REG_FILTER_MAX_DIST = 0, B200REG_FILTER_MIN_DIST = 1, B200REG_FILTER_BOUNDING_BOX = 2, B200REG_FILTER_RANDOM_SAMPLING = 3 };
typedef struct b200reg_filter_params {
  int32_t kind;
  int32_t dim;            /* MaxDist / MinDist: -1 radius, 0 x, 1 y, 2 z                         */
  float   dist;           /* maxDist | minDist                                                   */
  float   box[6];         /* xMin xMax yMin yMax zMin zMax                                       */
  int32_t remove_inside;
  int32_t random_method;  /* randomSamplingMethod: 0 direct RNG, 1 uniform_real_distribution     */
  double  prob;
  uint64_t seed;
} b200reg_filter_params;
int32_t b200reg_cloud_filter(b200reg_ctx* ctx, b200reg_cloud* inout, const b200reg_filter_params* params);
/* SurfaceNormalDataPointsFilter::inPlaceFilter (DataPointsFilters/SurfaceNormal.cpp:83-330).  Writes the cloud's
 * normals; optional HOST outputs (nullable): densities[n], eig_values[3n] ascending, lin_plan_curv[3n].
 * orient_origin (nullable float[3]): flip normals toward that point (CloudRegistration.cpp:71-74 usage). */
int32_t b200reg_surface_normals(b200reg_ctx* ctx, b200reg_cloud* inout, int32_t knn, float max_dist, float epsilon,
                                const float* orient_origin, float* densities_out, float* eig_values_out,
                                float* lin_plan_curv_out, int32_t* degenerate_count);

/* The complete descriptor set of the filter (SurfaceNormal.cpp:126-144,247-323).  Optional HOST outputs (nullable): densities[n],
 * eig_values[3n] ascending, eig_vectors[9n] (serializeEigVec: row-major, eigenvectors in columns, column 0 = the normal's), matched_ids
 * [knn n] (the neighbours' indices as floats, -1 = none; ascending (d2, index)), mean_dists[n] (|point - mean of its neighbours|;
 * numeric_limits<size_t>::max() for a degenerate point), lin_plan_curv[3n].  smooth_normals reproduces the reference's in-place
 * sweep exactly: point i averages the already smoothed normals of its lower-index neighbours and the original normals of the others,
 * each flipped to the side of its own original normal. */
typedef struct b200reg_normals_options {
  int32_t knn;
  float   max_dist, epsilon;
  int32_t smooth_normals;
  const float* orient_origin;   /* nullable float[3]: flip normals toward that point before smoothing */
  float *densities, *eig_values, *eig_vectors, *matched_ids, *mean_dists, *lin_plan_curv;
} b200reg_normals_options;
int32_t b200reg_surface_normals_ex(b200reg_ctx* ctx, b200reg_cloud* inout, const b200reg_normals_options* options, int32_t* degenerate_count);

/* ---- open3d_slam voxel-map crop / insert (fp64) --------------------------------------------------------------------- */
enum { B200REG_CROP_MAX_RADIUS = 0, B200REG_CROP_MIN_RADIUS = 1, B200REG_CROP_MINMAX_RADIUS = 2, B200REG_CROP_CYLINDER = 3 };
typedef struct b200reg_crop_spec {   /* CroppingVolume (open3d_slam/src/croppers.cpp:53-167) */
  int32_t kind;
  int32_t invert;
  double  center[3];   /* pose_.translation(); rotation ignored like the reference */
  double  rmin, rmax, zmin, zmax;
} b200reg_crop_spec;

int32_t b200reg_map_create(b200reg_ctx* ctx, double voxel_size, b200reg_map** out);
void    b200reg_map_destroy(b200reg_ctx* ctx, b200reg_map* map);
int64_t b200reg_map_size(const b200reg_map* map);
int32_t b200reg_map_has_normals(const b200reg_map* map);
/* Submap::insertScan (open3d_slam/src/Submap.cpp:39-96): transform (helpers.cpp:283-318), append, re-voxelise inside
 * `volume` (helpers.cpp:117-192).  scan_xyz / scan_nrm: HOST fp64 arrays (nrm nullable). */
int32_t b200reg_map_insert(b200reg_ctx* ctx, b200reg_map* map, const double* scan_xyz, const double* scan_nrm, int64_t n,
                           const double T_map_scan[16], const b200reg_crop_spec* volume);
/* Space carving on the resident map: Submap::carve (open3d_slam/src/Submap.cpp:130-144) -> getIdxsOfCarvedPoints
 * (helpers.cpp:245-281): the raw scan (HOST fp64, sensor frame) is moved into the map frame, every ray is sampled every
 * voxel_size from the sensor up to max(voxel_size, min(length - truncation_distance, max_raytracing_length)), and every map point
 * inside `volume` that shares a voxel with a sample is deleted unless the map has normals and |ray . normal| <=
 * min_dot_product_with_normal.  The surviving points keep their order.  removed (nullable): number of deleted points. */
typedef struct b200reg_carve_params {   /* SpaceCarvingParameters (Parameters.hpp:88-95) */
  double voxel_size, max_raytracing_length, truncation_distance, min_dot_product_with_normal;
} b200reg_carve_params;
int32_t b200reg_map_carve(b200reg_ctx* ctx, b200reg_map* map, const double* scan_xyz, int64_t n, const double T_map_scan[16],
                          const b200reg_crop_spec* volume, const b200reg_carve_params* params, int64_t* removed);
/* ScanToMapRegistration::cropSubmap (open3d_slam/src/ScanToMapRegistration.cpp:90-96) + open3dToPointmatcher:
 * order-preserving crop of the resident map into a float cloud ready for b200reg_icp_set_reference. */
int32_t b200reg_map_crop(b200reg_ctx* ctx, const b200reg_map* map, const b200reg_crop_spec* volume, b200reg_cloud** out);
int32_t b200reg_map_download(b200reg_ctx* ctx, const b200reg_map* map, double* xyz, double* nrm);
/* CroppingVolume::crop on a host fp64 cloud (croppers.cpp:76-106): returns the kept count, writes compacted output */
int32_t b200reg_crop_f64(b200reg_ctx* ctx, const b200reg_crop_spec* volume, const double* xyz, const double* nrm, int64_t n,
                         double* xyz_out, double* nrm_out, int64_t* n_out);

/* o3d_slam::voxelize (open3d_slam/src/helpers.cpp:108-115) -> Open3D v0.15.1 PointCloud::VoxelDownSample (external, pinned by
 * open3d_catkin/CMakeLists.txt:106-107) on a HOST fp64 cloud: per-voxel mean position and mean normal, grid anchored at
 * min_bound - voxel/2; output in ascending (x,y,z) key order.  xyz_out / nrm_out: room for n points. */
int32_t b200reg_voxel_down_sample_f64(b200reg_ctx* ctx, double voxel_size, const double* xyz, const double* nrm, int64_t n,
                                      double* xyz_out, double* nrm_out, int64_t* n_out);

/* ---- the scan-processing chain on device-resident fp64 clouds: ScanToMapIcp::processForScanMatchingAndMerging
 *      (open3d_slam/src/ScanToMapRegistration.cpp:36-69) without a host round trip between its steps.  Every `out` parameter is
 *      in/out: *out == NULL creates a cloud, otherwise the existing cloud's storage is reused (grown when too small). --------- */
int32_t b200reg_cloud64_upload(b200reg_ctx* ctx, const double* xyz, const double* nrm, int64_t n, b200reg_cloud64** out);
/* Next-scan prefetch (SlamWrapper's range-scan buffer feeding Mapper::addRangeMeasurement, SlamWrapper.cpp mapping worker): the
 * upload of b200reg_cloud64_upload enqueued on the context's copy stream, behind everything already enqueued on the compute
 * stream, so it overlaps the kernels of the scan being processed; returns without waiting (truly asynchronous for pinned host
 * memory).  The host buffers must stay unchanged until b200reg_cloud64_acquire, which waits for the copy and orders the compute
 * stream behind it; every other use of a prefetched cloud before acquire is undefined. */
int32_t b200reg_cloud64_prefetch(b200reg_ctx* ctx, const double* xyz, const double* nrm, int64_t n, b200reg_cloud64** out);
int32_t b200reg_cloud64_acquire(b200reg_ctx* ctx, b200reg_cloud64* cloud);
int32_t b200reg_cloud64_download(b200reg_ctx* ctx, const b200reg_cloud64* cloud, double* xyz, double* nrm);
int64_t b200reg_cloud64_size(const b200reg_cloud64* cloud);
int32_t b200reg_cloud64_has_normals(const b200reg_cloud64* cloud);
void    b200reg_cloud64_free(b200reg_ctx* ctx, b200reg_cloud64* cloud);
/* CroppingVolume::crop (croppers.cpp:76-106), order preserving */
int32_t b200reg_cloud64_crop(b200reg_ctx* ctx, const b200reg_crop_spec* volume, const b200reg_cloud64* in, b200reg_cloud64** out);
/* o3d_slam::voxelize (helpers.cpp:108-115), same convention as b200reg_voxel_down_sample_f64 */
int32_t b200reg_cloud64_voxel_down_sample(b200reg_ctx* ctx, double voxel_size, const b200reg_cloud64* in, b200reg_cloud64** out);
/* estimateNormalsOrCovariancesIfNeeded (open3d_slam/src/CloudRegistration.cpp:62-75) -> Open3D EstimateNormals(Hybrid(radius, knn)) +
 * OrientNormalsTowardsCameraLocation, in Open3D's scalar (fp64): the (up to) knn nearest points with d2 < radius^2, covariance and
 * eigenvector from the fp64 coordinates, flipped toward `origin` (nullable = no orientation).  The fp32 index only proposes neighbour
 * candidates (knn + 2, re-ranked in fp64).  Fewer than 3 neighbours: zero normal, counted in degenerate_count.  knn <= 32. */
int32_t b200reg_cloud64_estimate_normals(b200reg_ctx* ctx, b200reg_cloud64* inout, int32_t knn, double radius, const double origin[3],
                                         int32_t* degenerate_count);
/* open3dToPointmatcher (open3d_conversions.cpp:57-118): fp64 -> the fp32 DataPoints mirror (asynchronous on the ctx stream) */
int32_t b200reg_cloud64_to_f32(b200reg_ctx* ctx, const b200reg_cloud64* in, b200reg_cloud** out);
/* Submap::insertScan (Submap.cpp:39-96) with the scan already on the device */
int32_t b200reg_map_insert_cloud64(b200reg_ctx* ctx, b200reg_map* map, const b200reg_cloud64* scan, const double T_map_scan[16],
                                   const b200reg_crop_spec* volume);
/* b200reg_map_carve with the raw scan already on the device */
int32_t b200reg_map_carve_cloud64(b200reg_ctx* ctx, b200reg_map* map, const b200reg_cloud64* scan, const double T_map_scan[16],
                                  const b200reg_crop_spec* volume, const b200reg_carve_params* params, int64_t* removed);

/* ---- sensor_msgs/PointCloud2 bytes straight to device clouds: replaces
 *      RosPointCloud2Deserializer<float>::deserialize (pointmatcher_ros/src/RosPointCloud2Deserializer.cpp:9-38,213-232; features
 *      x y z pad = 1 and the "normals" descriptor) and open3d_conversions::rosToOpen3d (open3d_conversions.cpp:203-244).  Both read
 *      fields through PointCloud2ConstIterator<float>: value i = the 4 bytes at data + i * point_step + offset as a host-order float,
 *      whatever the declared datatype; row_step and is_bigendian are not consulted.  Offsets < 0 = field absent (normals are taken
 *      only when all three are present); x, y, z are required.  *out is in/out like the cloud64 entry points. ------------------- */
typedef struct b200reg_pc2_layout {
  uint32_t width, height, point_step;
  int32_t off_x, off_y, off_z;
  int32_t off_normal_x, off_normal_y, off_normal_z;
} b200reg_pc2_layout;
int32_t b200reg_cloud_from_pointcloud2(b200reg_ctx* ctx, const uint8_t* data, size_t data_bytes, const b200reg_pc2_layout* layout,
                                       b200reg_cloud** out);
int32_t b200reg_cloud64_from_pointcloud2(b200reg_ctx* ctx, const uint8_t* data, size_t data_bytes, const b200reg_pc2_layout* layout,
                                         b200reg_cloud64** out);

/* ---- loop-closure support operations around b200reg_icp_register_batch --------------------------------- */
/* computeIndicesOfOverlappingPoints (open3d_slam/src/helpers.cpp:320-345; call site PlaceRecognition.cpp:100-106): source is moved
 * by T (PointCloud::Transform), both clouds are binned with key = floor(p / voxel_size); a point is in the overlap when its voxel
 * holds >= min_points_per_voxel source AND target points.  idx_*_out (HOST, room for the cloud's size, nullable): ascending point
 * indices (the reference emits them in unordered_map order — a set either way). */
int32_t b200reg_overlap_indices(b200reg_ctx* ctx, const b200reg_cloud64* source, const b200reg_cloud64* target,
                                const double T_source_to_target[16], double voxel_size, int64_t min_points_per_voxel,
                                int64_t* idx_source_out, int64_t* n_source_out, int64_t* idx_target_out, int64_t* n_target_out);
/* open3d::pipelines::registration::GetInformationMatrixFromPointClouds (Open3D v0.15.1, external; call sites
 * PlaceRecognition.cpp:148-149, constraint_builders.cpp:71-72): every source point moved by T takes its nearest target point
 * within max_correspondence_distance (the path's fp32 k = 1 search); info_out = sum over correspondences of G^T G with
 * G = [-[q]x | I] of the matched TARGET point q, 6x6 row-major, unknown order (rotation, translation), accumulated in fp64. */
int32_t b200reg_information_matrix(b200reg_ctx* ctx, const b200reg_cloud64* source, const b200reg_cloud64* target,
                                   double max_correspondence_distance, const double T_source_to_target[16], double info_out[36],
                                   int64_t* n_correspondences);

#ifdef __cplusplus
}
#endif
#endif /* B200REG_H */
