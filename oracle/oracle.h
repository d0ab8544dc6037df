/*
 * oracle/oracle.h — TEST INFRASTRUCTURE ONLY (never linked into, imported by or executed from the product).
 *
 * C API of the CPU oracle: a dependency-free restatement of the reference's scan-to-map registration hot path
 * (libpointmatcher ICP chain + open3d_slam voxel-map crop/insert).  The reference itself cannot be compiled in
 * this environment (Eigen, boost, yaml-cpp, libnabo, Open3D are absent), so each function restates the algorithm of
 * the reference file:line cited beside its definition in oracle.cpp.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load liboracle.so.
 *
 * Layout conventions (identical to the reference's Eigen column-major matrices):
 *   features : `stride` floats per point, xyz first (stride 4 = x y z pad, pad == 1)
 *   normals  : 3 floats per point
 *   matches  : ids[k*i + j], d2[k*i + j] = j-th neighbour of reading point i (k contiguous values per point),
 *              invalid = (-1, +inf); distances are SQUARED (PointMatcher.h:444-451)
 *   4x4 transforms are ROW-MAJOR float[16] at this C boundary.
 */
#ifndef B200REG_ORACLE_H
#define B200REG_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* precision of the error-minimiser arithmetic */
enum { ORC_REF32 = 0 /* fp32 everywhere the reference is fp32 */, ORC_MIXED = 1 /* fp32 data path, fp64 sums+solve (the GPU contract) */ };

typedef struct orc_icp_config {
  int32_t matcher_kind;   /* 0 = KDTreeMatcher, 1 = MirrorMatcher (fixture for Conditioning tests) */
  int32_t knn;            /* MatchersImpl.h:97 */
  float   max_dist;       /* MatchersImpl.h:100 (inf allowed) */
  float   epsilon;        /* parsed, treated as 0 */
  float   trim_ratio;     /* TrimmedDistOutlierFilter ratio; <= 0 : filter absent */
  float   max_angle;      /* SurfaceNormalOutlierFilter maxAngle; < 0 : filter absent */
  int32_t force2D;        /* PointToPlane.h:74 */
  int32_t force4DOF;      /* PointToPlane.h:75 */
  int32_t max_iter;       /* CounterTransformationChecker; < 0 : absent (0 stops after the first iteration, like 1) */
  float   min_diff_rot;   /* DifferentialTransformationChecker */
  float   min_diff_trans;
  int32_t smooth_length;  /* < 0 : checker absent */
  float   bound_rot;      /* BoundTransformationChecker; < 0 : absent */
  float   bound_trans;
  int32_t precision;      /* ORC_REF32 | ORC_MIXED */
  int32_t nthreads;       /* OpenMP threads for kNN (1 = the reference's execution model) */
} orc_icp_config;

typedef struct orc_icp_stats {
  int32_t iterations;               /* IterationsCount (ICP.cpp:452) */
  int32_t max_iterations_reached;   /* ICP.cpp:444 */
  int32_t status;                   /* 0 ok, <0 error (see ORC_E_*) */
  int32_t kept_last;                /* pairs kept in the last iteration */
  uint64_t points_touched;          /* PointCountTouched (ICP.cpp:453) */
  float   point_used_ratio;         /* ErrorMinimizer.cpp:139 */
  float   weighted_point_used_ratio;/* ErrorMinimizer.cpp:140 = OverlapRatio stat */
  float   last_tau;                 /* last trimmed-distance limit */
  float   mean_ref[3];              /* T_refIn_refMean translation */
  double  t_reference_s;            /* ReferencePreprocessingDuration */
  double  t_converge_s;             /* ConvergenceDuration */
} orc_icp_stats;

enum { ORC_OK = 0, ORC_E_CONVERGENCE = -1, ORC_E_TRANSFORM = -2, ORC_E_PARAM = -3, ORC_E_EMPTY = -4, ORC_E_FIELD = -5 };

/* ---- matcher (MatchersImpl.cpp:109-132 + libnabo public contract) ---- */
void*    orc_kdtree_build(const float* feat, int32_t stride, int64_t n);
void     orc_kdtree_free(void* tree);
uint64_t orc_kdtree_knn(const void* tree, const float* q, int32_t qstride, int64_t nq, int32_t k, float max_dist,
                        int32_t allow_self, int32_t* ids, float* d2, int32_t nthreads);
uint64_t orc_knn_brute(const float* ref, int32_t rstride, int64_t m, const float* q, int32_t qstride, int64_t nq, int32_t k,
                       float max_dist, int32_t* ids, float* d2, int32_t nthreads);

/* ---- RigidTransformation (TransformationsImpl.cpp:61-114) ---- */
int32_t orc_transform(const float T[16], float* feat, int32_t stride, float* normals, int64_t n);

/* ---- outlier filters ---- */
int32_t orc_dist_outlier(int32_t kind, const float* d2, int64_t count, float param, float* limit_out, float* w_out);   /* OutlierFiltersImpl.cpp:64-125 */
int32_t orc_trimmed_dist(const float* d2, int64_t count, float ratio, float* tau_out, float* w_out);   /* OutlierFiltersImpl.cpp:140-147, Matches.cpp:61-87 */
void    orc_surface_normal_outlier(const float* nrm_read, const float* nrm_ref, const int32_t* ids, int32_t k, int64_t n,
                                   float max_angle, float* w_out);                                     /* OutlierFiltersImpl.cpp:226-281 */
void    orc_no_filter_weights(const float* d2, int64_t count, float* w_out);                            /* OutlierFilter.cpp:70-85 */

/* ---- ErrorElements + PointToPlaneErrorMinimizer (ErrorMinimizer.cpp:59-193, PointToPlane.cpp:107-368) ----
 * A_out (dim x dim row-major, dim = 6 | 4 (force4DOF) | 3 (force2D)), b_out, x_out are filled for inspection. */
int32_t orc_point_to_plane(const float* read_feat, int32_t rstride, int64_t n, const float* ref_feat, int32_t fstride,
                           const float* ref_normals, const int32_t* ids, const float* d2, const float* w, int32_t k,
                           int32_t force2D, int32_t force4DOF, int32_t precision, float T_out[16], double* A_out, double* b_out,
                           double* x_out, int32_t* kept_out, float* point_used_ratio, float* weighted_point_used_ratio);

/* ---- whole registration (ICP.cpp:258-468) ---- */
int32_t orc_icp(const orc_icp_config* cfg, const float* ref_feat, int32_t ref_stride, const float* ref_normals, int64_t m,
                const float* read_feat, int32_t read_stride, const float* read_normals, int64_t n, const float T_init[16],
                float T_out[16], orc_icp_stats* stats);

/* the same, split the way Mapper uses it (Mapper.cpp:363,391-393): initReference once, compute many times */
void*   orc_icp_ref_create(const orc_icp_config* cfg, const float* ref_feat, int32_t ref_stride, const float* ref_normals, int64_t m);
void    orc_icp_ref_free(void* ref);
int32_t orc_icp_compute(const orc_icp_config* cfg, const void* ref, const float* read_feat, int32_t read_stride,
                        const float* read_normals, int64_t n, const float T_init[16], float T_out[16], orc_icp_stats* stats);

/* ---- VoxelGridDataPointsFilter (VoxelGrid.cpp:71-344); in place, returns the number of points kept or <0 ---- */
int64_t orc_voxelgrid(float* feat, int32_t stride, float* desc, int32_t desc_dim, int64_t n, float vx, float vy, float vz,
                      int32_t use_centroid, int32_t average_descriptors, uint32_t* cell_of_point_out /* nullable, n entries */);

/* ---- SurfaceNormalDataPointsFilter (SurfaceNormal.cpp:83-330); outputs nullable ---- */
int32_t orc_surface_normals(const float* feat, int32_t stride, int64_t n, int32_t knn, float max_dist, int32_t precision,
                            int32_t nthreads, const float* orient_origin /* nullable: flip toward this point */,
                            float* normals, float* densities, float* eig_values, float* eig_vectors, float* matched_ids,
                            float* mean_dists, float* lin_plan_curv, int32_t* degenerate_count);

/* PointToPlaneErrorMinimizer::getResidualError (PointToPlane.cpp:372-422) */
int32_t orc_point_to_plane_residual(const float* read_feat, int32_t rstride, int64_t n, const float* ref_feat, int32_t fstride,
                                    const float* ref_normals, const int32_t* ids, const float* d2, const float* w, int32_t k,
                                    int32_t force2D, int32_t precision, double* residual_out);
/* smoothNormals (SurfaceNormal.cpp:299-323): the in-place sequential sweep, on the normals orc_surface_normals returned */
int32_t orc_smooth_normals(const float* feat, int32_t stride, int64_t n, int32_t knn, float max_dist, int32_t nthreads, float* normals);

/* Open3D EstimateNormals(Hybrid(radius, knn)) + orientation toward `origin` (CloudRegistration.cpp:62-75), fp64 throughout */
int32_t orc_estimate_normals_f64(const double* xyz, int64_t n, int32_t knn, double radius, const double* origin /* nullable */, int32_t exact,
                                 int32_t nthreads, double* normals, int32_t* degenerate_count);

/* ---- open3d_slam side (fp64) ---- */
enum { ORC_CROP_MAX_RADIUS = 0, ORC_CROP_MIN_RADIUS = 1, ORC_CROP_MINMAX_RADIUS = 2, ORC_CROP_CYLINDER = 3 };
typedef struct orc_crop_spec {
  int32_t kind; int32_t invert;
  double center[3];            /* pose_.translation() — rotation is ignored (croppers.cpp:122) */
  double rmin, rmax, zmin, zmax;
} orc_crop_spec;
int64_t orc_crop(const orc_crop_spec* spec, const double* xyz, const double* nrm, int64_t n, double* xyz_out, double* nrm_out,
                 uint8_t* mask_out);                                                  /* croppers.cpp:53-167 */
void    orc_transform_f64(const double T[16], const double* xyz, const double* nrm, int64_t n, double* xyz_out, double* nrm_out); /* helpers.cpp:283-318 (no identity-duplication) */
int64_t orc_voxelize_within_volume(double voxel, const orc_crop_spec* spec, const double* xyz, const double* nrm, int64_t n,
                                   double* xyz_out, double* nrm_out, int64_t* n_passthrough);   /* helpers.cpp:117-192; output = pass-through, then voxels in ascending (x,y,z) key order */
int64_t orc_voxel_down_sample(double voxel, const double* xyz, const double* nrm, int64_t n, double* xyz_out, double* nrm_out); /* Open3D v0.15.1 VoxelDownSample, ascending key order */
int64_t orc_carve(const double* scan_xyz, int64_t n_scan, const double sensor[3], const double* map_xyz, const double* map_nrm,
                  int64_t n_map, const orc_crop_spec* spec, double voxel, double max_ray_length, double truncation, double min_dot,
                  uint8_t* remove_mask);   /* Submap.cpp:130-144 + helpers.cpp:245-281; scan already in the map frame */
int32_t orc_pointcloud2(const uint8_t* data, int64_t n, uint32_t point_step, const int32_t offs[6], float* feat4, float* nrm3,
                        double* xyz64, double* nrm64);   /* RosPointCloud2Deserializer.cpp:9-38,113-143,213-232; open3d_conversions.cpp:203-244 */
void    orc_overlap(const double* src, int64_t n_src, const double* tgt, int64_t n_tgt, const double T[16], double voxel, int64_t min_pts,
                    uint8_t* src_mask, uint8_t* tgt_mask);   /* helpers.cpp:320-345 */
int64_t orc_information_matrix(const double* src, int64_t n_src, const double* tgt, int64_t n_tgt, double max_dist, const double T[16],
                               double info[36]);   /* Open3D v0.15.1 GetInformationMatrixFromPointClouds (external) */
void    orc_o3d_to_pm(const double* xyz, const double* nrm, int64_t n, float* feat4, float* nrm3);   /* open3d_conversions.cpp:57-118 */

/* ---- SURVEY 8f rank 3 modules ---- */
int32_t orc_var_trimmed_dist(const float* d2, int64_t count, float min_ratio, float max_ratio, float lambda, float* ratio_out, float* tau_out,
                             float* w_out);   /* OutlierFiltersImpl.cpp:152-220 */
int32_t orc_point_to_point(const float* read_feat, int32_t rstride, int64_t n, const float* ref_feat, int32_t fstride, const int32_t* ids,
                           const float* d2, const float* w, int32_t k, int32_t precision, float T_out[16],
                           int32_t* kept_out);   /* ErrorMinimizers/PointToPoint.cpp:55-94 */
int32_t orc_point_to_plane_covariance(const float* read_feat, int32_t rstride, int64_t n, const float* ref_feat, int32_t fstride,
                                      const float* ref_normals, const int32_t* ids, const float* d2, const float* w, int32_t k,
                                      int32_t precision, const float T[16], float sensor_std_dev,
                                      double cov_out[36]);   /* ErrorMinimizers/PointToPlaneWithCov.cpp:70-160 */
int64_t orc_cloud_filter(const float* feat, int32_t stride, int64_t n, int32_t kind, int32_t dim, float dist, const float box[6],
                         int32_t remove_inside, int32_t random_method, double prob, uint64_t seed,
                         uint8_t* keep);   /* DataPointsFilters/MaxDist.cpp, MinDist.cpp, BoundingBox.cpp, RandomSampling.cpp */

#ifdef __cplusplus
}
#endif
#endif
