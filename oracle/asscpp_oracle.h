/*
 * asscpp_oracle.h -- CPU restatement ("oracle") of the kucars/asscpp viewpoint
 * coverage-evaluation hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (asscpp_b200/) never
 * links, imports or calls it.
 *
 * PARITY UNPINNED: the reference ships no tests, golden vectors or recorded
 * outputs for this path and cannot be compiled here
 * (PCL / Eigen / tf / ROS are absent), so the restatement is anchored on the
 * cited source lines plus the sensor-pose fixtures under cscpp/txt.  The
 * arithmetic that lives in un-vendored PCL / Eigen / tf (versions unpinned by
 * the reference) is restated from their published algorithms with the
 * canonical float32 operation orders listed in SURVEY.md section 8c.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/cscpp unless prefixed with sspp/).
 */
#ifndef ASSCPP_ORACLE_H_
#define ASSCPP_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OrcParams {
  float leaf;          /* occlusion voxel size; src/occlusion_culling.cpp:41 (0.5f)          */
  float hfov_deg;      /* src/occlusion_culling.cpp:74 (58)                                  */
  float vfov_deg;      /* src/occlusion_culling.cpp:73 (45)                                  */
  float near_dist;     /* src/occlusion_culling.cpp:75 (0.7)                                 */
  float far_dist;      /* src/occlusion_culling.cpp:76 (6.0)                                 */
  float pcd_origin[3]; /* model cloud sensor_origin_ (PCD VIEWPOINT), occlusion_culling.cpp:338 */
  float cov_leaf;      /* coverage lattice, coverage_path_planning_heuristic.cpp:479 (0.5f)  */
} OrcParams;

/* per-view statistics slots of orc_eval_view / orc_eval_batch */
enum {
  ORC_ST_SURVIVORS = 0, /* S_v: frustum survivors                                   */
  ORC_ST_VISIBLE = 1,   /* Vis_v: points returned by extractVisibleSurface          */
  ORC_ST_CARRY = 2,     /* Q5 carry events (occlusionEstimation returned -1)        */
  ORC_ST_UNIQUE = 3,    /* U_v: occupied cells of the per-view grid                 */
  ORC_ST_CELLS = 4,     /* G_v: cells of the per-view grid (div_x*div_y*div_z)      */
  ORC_ST_STEPS = 5,     /* DDA loop iterations executed (memoised: per unique ray)  */
  ORC_ST_GAIN_PTS = 6,  /* |visible \ covered_points|                               */
  ORC_ST_GAIN_VOX = 7,  /* |cov-voxels(visible) \ covered_voxels|                   */
  ORC_ST_HASH = 8,      /* order-independent 64-bit hash of the visible index set   */
  ORC_ST_SURV_HASH = 9, /* same for the survivor index set                          */
  ORC_ST_COUNT = 10
};

/* epsilon classes of SURVEY.md 8c, counted when requested */
enum { ORC_EPS_E1 = 0, ORC_EPS_E2 = 1, ORC_EPS_E3 = 2, ORC_EPS_E4 = 3, ORC_EPS_E5 = 4, ORC_EPS_COUNT = 5 };

typedef struct OrcCtx OrcCtx;

int orc_create(const float* xyz, int64_t n, const OrcParams* p, OrcCtx** out);
void orc_destroy(OrcCtx* c);

/* model-level facts */
int64_t orc_num_points(const OrcCtx* c);
int64_t orc_num_cov_voxels(const OrcCtx* c);   /* OriginalVoxelsSize, occlusion_culling.cpp:50-65 */
int64_t orc_num_occ_voxels(const OrcCtx* c);   /* occupied voxels on the occlusion lattice        */
void orc_model_box(const OrcCtx* c, int32_t min_b[3], int32_t max_b[3]); /* occlusion lattice */
/* compact id (0..num_cov_voxels) of the coverage voxel of every point, ids ascending in
 * (k, j, i) order of the model grid */
void orc_point_cov_ids(const OrcCtx* c, int32_t* out);

/* frustum planes of a pose (x y z qx qy qz qw); planes[6][4] in order l r t b f n; T[3]
 * frustum_culling.cpp:75-141, occlusion_culling.cpp:271-290 */
void orc_planes(const OrcParams* p, const double pose[7], float planes[24], float T[3]);
/* 8 frustum corners: far tl tr bl br, near tl tr bl br (frustum_culling.cpp:97-107) */
void orc_corners(const OrcParams* p, const double pose[7], float corners[24]);

/* One viewpoint, literal or memoised.  surv_out / vis_out may be NULL; otherwise they need
 * room for n ints.  covered_pts: bitset over point index (bit i of word i/64), or NULL;
 * covered_vox: bitset over compact coverage-voxel id, or NULL.  stats: ORC_ST_COUNT slots.
 * eps: NULL or ORC_EPS_COUNT slots (accumulated).  memo != 0 evaluates each occupied cell
 * once (Q1) -- results are identical by construction and tests check it. */
int orc_eval_view(OrcCtx* c, const double pose[7], int memo, const uint64_t* covered_pts,
                  const uint64_t* covered_vox, int32_t* surv_out, int32_t* vis_out,
                  int64_t stats[ORC_ST_COUNT], int64_t* eps);

/* V viewpoints on nthreads host threads (each view is independent). stats: V*ORC_ST_COUNT. */
int orc_eval_batch(OrcCtx* c, const double* poses, int64_t nviews, int memo, int nthreads,
                   const uint64_t* covered_pts, const uint64_t* covered_vox, int64_t* stats);

/* accumulate the visible sets of the given views into caller-owned bitsets (set union,
 * coverage_path_planning_heuristic.cpp:344-346): only views with > min_visible visible points. */
int orc_accumulate(OrcCtx* c, const double* poses, int64_t nviews, int nthreads,
                   int64_t min_visible, uint64_t* covered_pts, uint64_t* covered_vox);

/* calcCoveragePercent of a point-index set (occlusion_culling.cpp:375-444), float32 */
float orc_coverage_percent_idx(const OrcCtx* c, const int32_t* idx, int64_t n);
/* same from a coverage-voxel bitset */
float orc_coverage_percent_vox(const OrcCtx* c, const uint64_t* covered_vox);
int64_t orc_popcount(const uint64_t* bits, int64_t nbits);

/* sspp/src/sensors.cpp:47-90 */
void orc_robot2sensor(const double robot_pose[7], const double offset[3], const double rpy[3],
                      double sensor_pose[7]);

/* tf::Matrix3x3(q).getRPY(roll, pitch, yaw) as called at occlusion_culling.cpp:497-500 / :558-561 */
void orc_pose_rpy(const double pose[7], double rpy[3]);
/* OcclusionCulling::pointCloudViewportTransform, occlusion_culling.cpp:538-618 (n x xyz float32 -> same) */
void orc_viewport_transform(const double pose[7], const float* xyz, int64_t n, float* out);
/* calcAvgAccuracy(cloud, pose) :459-475 (err_sum_depth / n) and calcAvgAccuracy(cloud) :445-458 (err_sum_z / n)
 * over the model points idx[0..n) in that order, plus the depth extrema of SSMaxMinAccuracy :505-513 */
void orc_accuracy_idx(const OrcCtx* c, const double pose[7], const int32_t* idx, int64_t n, double* err_sum_depth,
                      double* err_sum_z, float* depth_min, float* depth_max);

/* pointCloudDiff (coverage_path_planning_heuristic.cpp:88-145) in bitset form: uncovered coverage voxels in the
 * reference's k, j, i scan order (= ascending compact id) with their cell centres; returns their number */
int64_t orc_uncovered(const OrcCtx* c, const uint64_t* covered_vox, int32_t* ids, float* centres);
/* pcl_aliasing != 0: with the reference's unchecked linear-index lookup (quirk Q11, heuristic.cpp:120-123) */
int64_t orc_uncovered_ex(const OrcCtx* c, const uint64_t* covered_vox, int pcl_aliasing, int32_t* ids, float* centres);
/* clusteringPointCloud (:147-190), <PCL EuclideanClusterExtraction>: labels[i] = cluster rank (largest first) or -1;
 * ties (optional) counts point pairs at squared distance exactly float(tol*tol) */
int32_t orc_euclidean_clusters(const float* xyz, int64_t n, double tol, int32_t min_size, int32_t max_size,
                               int32_t* labels, int64_t* ties);
/* findClusterBB (:218-246) */
void orc_cluster_bb(const float* xyz, int64_t n, float leaf, double grid_start[3], double grid_size[3]);

/* mesh filters of the adaptive loop (coverage_path_planning_heuristic.cpp:300-335); CGAL restated, see the .cpp */
typedef struct OrcMesh OrcMesh;
OrcMesh* orc_mesh_create(const float* vertices, int64_t nv, const int32_t* triangles, int64_t nt);
void orc_mesh_destroy(OrcMesh* m);
int64_t orc_mesh_num_triangles(const OrcMesh* m);
/* graze (optional): smallest normalised distance of the ray's crossing parameters from a triangle border */
void orc_mesh_filter(const OrcMesh* m, const double* pos, int64_t n, const double through[3], double min_dist,
                     double max_dist, int nthreads, int32_t* hits, double* sqd, uint8_t* pass, double* graze);
void orc_mesh_segments_free(const OrcMesh* m, const double* a, const double* b, int64_t n, int nthreads, uint8_t* out);

/* 64-bit mix used for the order-independent set hashes (sum of mix(idx)) */
uint64_t orc_mix64(uint64_t x);

#ifdef __cplusplus
}
#endif
#endif
