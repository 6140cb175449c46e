/*
 * asscpp_gpu.h -- C ABI of the B200-native coverage-evaluation path.
 *
 * This is the drop-in boundary.  The reference (kucars/asscpp) has no FFI of its own: its hot
 * path is the C++ class OcclusionCulling (cscpp/include/cscpp/occlusion_culling.h:50-95) called
 * by SSPP::CoveragePathPlanningHeuristic (cscpp/src/coverage_path_planning_heuristic.cpp).  The
 * functions below are what a C++ shim of that class (include/cscpp/occlusion_culling_b200.h) and
 * the viewpoint batcher bind to.  Each entry point names the reference code it replaces.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a negative
 * ASSCPP_E* code; no exceptions cross the boundary; one context per GPU; calls on one context
 * must come from one host thread at a time.  There is no CPU fallback: without a CUDA device
 * asscpp_create fails with ASSCPP_ENODEVICE.
 */
#ifndef ASSCPP_GPU_H_
#define ASSCPP_GPU_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ASSCPP_ABI_VERSION 1

enum {
  ASSCPP_OK = 0,
  ASSCPP_EINVAL = -1,     /* bad argument                                                  */
  ASSCPP_ENODEVICE = -2,  /* no usable CUDA device (there is no CPU path)                   */
  ASSCPP_ECUDA = -3,      /* CUDA runtime error, text in asscpp_last_error                  */
  ASSCPP_ERANGE = -4,     /* grid too large (PCL: "Leaf size is too small") / coords not finite */
  ASSCPP_ENOMEM = -5,
  ASSCPP_ESTATE = -6,     /* call order (e.g. results requested before an eval)             */
  ASSCPP_ENCCL = -7       /* NCCL unavailable or failed                                     */
};

/* Hard-coded literals of the reference turned into parameters.
 * Defaults: occlusion_culling.cpp:41 (voxelRes 0.5f), :73-76 (VFOV 45, HFOV 58, near 0.7, far 6.0),
 * pcd_origin = PCD VIEWPOINT translation used at occlusion_culling.cpp:338-341,
 * cov_leaf = coverage_path_planning_heuristic.cpp:479 (0.5f). */
typedef struct AsscppParams {
  float leaf;
  float hfov_deg;
  float vfov_deg;
  float near_dist;
  float far_dist;
  float pcd_origin[3];
  float cov_leaf;
} AsscppParams;

typedef struct AsscppModelInfo {
  int64_t n_points;
  int64_t n_occ_voxels;  /* occupied voxels, occlusion lattice                        */
  int64_t n_cov_voxels;  /* OriginalVoxelsSize (occlusion_culling.cpp:50-65)          */
  int64_t n_chunks;      /* leaf buckets of the cull hierarchy                        */
  int32_t min_b[3];      /* model box on the occlusion lattice                        */
  int32_t max_b[3];
  int32_t sm_count;
  int32_t reserved;
} AsscppModelInfo;

/* Per-viewpoint result of asscpp_eval_batch (one "VE"). */
typedef struct AsscppViewResult {
  uint32_t survivors;   /* S_v  : points passing FrustumCullingTT::applyFilter (frustum_culling.cpp:150-171) */
  uint32_t visible;     /* Vis_v: size of extractVisibleSurface's cloud (occlusion_culling.cpp:371)          */
  uint32_t carry;       /* Q5 events: occlusionEstimation returned -1 (occlusion_culling.cpp:332)             */
  uint32_t unique_rays; /* U_v  : occupied cells of the per-view grid = distinct DDA rays                     */
  uint32_t gain_pts;    /* |visible \ covered points|                                                         */
  uint32_t gain_vox;    /* |coverage voxels(visible) \ covered voxels| (heuristic.cpp:473-483 in bitset form) */
  uint32_t status;      /* ASSCPP_VS_* bits                                                                   */
  uint32_t reserved;
  uint64_t cells;       /* G_v: cells of the per-view grid                                                    */
  uint64_t steps;       /* DDA loop iterations over the unique rays (early exit at the first occluder)        */
  uint64_t vis_hash;    /* sum of mix64(original index) over the visible set   (ASSCPP_EVAL_HASH)             */
  uint64_t surv_hash;   /* same over the survivors                              (ASSCPP_EVAL_HASH)            */
} AsscppViewResult;

enum {
  ASSCPP_VS_GRID_SHARED = 1u, /* per-view bit-grid lived in shared memory                    */
  ASSCPP_VS_GRID_GLOBAL = 2u, /* ... in the per-CTA global (L2-resident) scratch             */
  ASSCPP_VS_ORDERED = 4u      /* a Q5 carry happened; view re-evaluated by the ordered path  */
};

/* flags of asscpp_eval_batch */
enum {
  ASSCPP_EVAL_HASH = 1u,        /* also compute vis_hash / surv_hash                               */
  ASSCPP_EVAL_BRUTE = 2u,       /* no bucket pre-cull: stream every point for every view           */
  ASSCPP_EVAL_FORCE_GLOBAL = 4u, /* keep every per-view grid in global memory (test / profiling)   */
  ASSCPP_EVAL_ACCURACY = 8u,     /* also reduce the accuracy terms of the visible set (asscpp_get_accuracy) */
  ASSCPP_EVAL_VIEW_CELLS = 16u   /* also keep, per view, the coverage voxels its visible cloud touches
                                    (asscpp_get_view_cells); at most 4 096 views per call, groups of at most 8,
                                    no accumulator set */
};

/* Accuracy terms of one viewpoint's visible cloud (ASSCPP_EVAL_ACCURACY).
 *   err_sum_depth / visible = OcclusionCulling::calcAvgAccuracy(cloud, pose)  (occlusion_culling.cpp:459-475)
 *   err_sum_z     / visible = OcclusionCulling::calcAvgAccuracy(cloud)        (occlusion_culling.cpp:445-458)
 *   depth_min / depth_max   = what SSMaxMinAccuracy folds over the search space (occlusion_culling.cpp:505-513)
 * depth = x of pointCloudViewportTransform (:538-618), float32.  The sums are the EXACT sums of the per-point
 * double terms 0.0000285*val*val, rounded once; the reference's left-to-right double sum differs from that by at
 * most visible/2 ulp. */
typedef struct AsscppViewAccuracy {
  double err_sum_depth;
  double err_sum_z;
  float depth_min;   /* +inf when nothing is visible */
  float depth_max;   /* -inf when nothing is visible */
  uint32_t visible;
  uint32_t reserved;
} AsscppViewAccuracy;

typedef struct AsscppCtx AsscppCtx;

/* -- lifetime ------------------------------------------------------------------------------- */
/* Replaces OcclusionCulling::OcclusionCulling (occlusion_culling.cpp:26-79): takes the model
 * cloud (n x {x,y,z} float32, host memory), builds the device-resident point SoA, the model
 * voxel tables and OriginalVoxelsSize.  device = CUDA ordinal. */
int asscpp_create(int device, const float* xyz, int64_t n, const AsscppParams* params, AsscppCtx** out);
int asscpp_destroy(AsscppCtx* ctx);
const char* asscpp_last_error(const AsscppCtx* ctx); /* ctx may be NULL: last create error */
int asscpp_abi_version(void);
int asscpp_model_info(const AsscppCtx* ctx, AsscppModelInfo* out);
void asscpp_default_params(AsscppParams* p);

/* -- batched viewpoint evaluation ------------------------------------------------------------ */
/* Replaces the per-viewpoint call OcclusionCulling::extractVisibleSurface (occlusion_culling.cpp:
 * 265-373) as issued by calculateHeuristic (heuristic.cpp:439-445) and isFilteringConditionSatisfied
 * (:339-354), for nviews poses at once.  poses: nviews x {x y z qx qy qz qw} doubles, host memory.
 * covered_set: handle from asscpp_covered_new to score gains against, or -1 for the empty set.
 * Asynchronous on the context's stream; results are fetched with asscpp_get_results. */
int asscpp_eval_batch(AsscppCtx* ctx, const double* poses, int64_t nviews, int covered_set, uint32_t flags);
/* Child scoring for one A* expansion (sspp/src/astar.cpp:311-330): views [g * group_size, (g + 1) * group_size)
 * are the sensor poses of candidate g (Node::senPoses).  Same evaluation as asscpp_eval_batch, except that
 * gain_vox of a view counts only coverage voxels that are neither in covered_set nor seen by an earlier view
 * of its group, so the sum over a group is the gain of the union of the sensors' visible clouds
 * (calculateHeuristic's visibleCloud, heuristic.cpp:439-445, scored as :473-483).  gain_pts stays per view.
 * Batches of up to 4096 views run in split-group mode: every VIEW is a work unit of its own (a group per CTA would
 * leave most SMs idle for the few dozen children of an A* expansion), each writes the coverage-voxel bitset of its
 * visible cloud, and a second small kernel derives the per-view gains; the numbers are the same.  In a larger batch a
 * group containing a view with ASSCPP_VS_ORDERED set (quirk Q5 replay) is not union-exact: score that group with
 * asscpp_covered_clone + asscpp_covered_or_views instead (split-group mode is exact there too). */
int asscpp_eval_groups(AsscppCtx* ctx, const double* poses, int64_t nviews, int group_size, int covered_set,
                       uint32_t flags);
int asscpp_get_results(AsscppCtx* ctx, AsscppViewResult* out, int64_t nviews); /* synchronises */
int asscpp_get_counts(AsscppCtx* ctx, uint32_t* survivors, uint32_t* visible, uint32_t* carry, int64_t nviews);
/* The two scores a search loop reads per view (c = coverage - parent coverage, heuristic.cpp:482-483, in bitset
 * form).  Moves 8 bytes per view device -> host, not the 64-byte record.  Either output may be NULL. */
int asscpp_get_gain(AsscppCtx* ctx, uint32_t* gain_pts, uint32_t* gain_vox, int64_t nviews);

/* General form of the evaluation calls above.  One launch does, per view: cull + per-view grid + DDA, gains against
 * covered_set (-1: the empty set), and -- for views whose visible count exceeds accum_min_visible -- the OR of the
 * visible set into accum_set (-1: none; must differ from covered_set).  That is isFilteringConditionSatisfied's
 * "if(pts.size()>5) globalCloud += pts" (heuristic.cpp:342-346) and calculateHeuristic's scoring (:473-483) in one
 * pass.  group_size as asscpp_eval_groups (0 or 1: every view on its own).  poses_on_device != 0: poses is a device
 * pointer (nviews x 7 doubles already in HBM); ASSCPP_EVAL_ACCURACY then uses device trigonometry. */
typedef struct AsscppEvalOpts {
  int32_t covered_set;
  int32_t accum_set;
  int64_t accum_min_visible;
  int32_t group_size;
  uint32_t flags;
  uint32_t poses_on_device;
  uint32_t reserved;
} AsscppEvalOpts;
int asscpp_eval_ex(AsscppCtx* ctx, const double* poses, int64_t nviews, const AsscppEvalOpts* opts);

/* Per-view coverage-voxel sets of the last evaluation issued with ASSCPP_EVAL_VIEW_CELLS: nviews rows of
 * asscpp_view_cells_words(ctx) 32-bit words, bit i of a row = the view's visible cloud has a point in the coverage voxel
 * with compact id i (the ids of asscpp_uncovered_voxels).  This is what the per-voxel heuristics of the reference walk:
 * the occupied octomap keys of a sensor's visible cloud (coverage_path_planning_heuristic.cpp:657-674 -- at the
 * reference's volumetricVoxelRes = 0.5 m the octomap cell of a point is its coverage voxel).  Synchronises. */
int64_t asscpp_view_cells_words(const AsscppCtx* ctx);
int asscpp_get_view_cells(AsscppCtx* ctx, uint32_t* rows, int64_t nviews);
/* accuracy terms of the last evaluation issued with ASSCPP_EVAL_ACCURACY (synchronises) */
int asscpp_get_accuracy(AsscppCtx* ctx, AsscppViewAccuracy* out, int64_t nviews);

/* Same evaluation with device-resident poses (nviews x 7 doubles already in HBM) and results left
 * on the device (AsscppViewResult[nviews], device pointer returned).  Used by the batcher when
 * poses are generated on the device and by the kernel-only benchmark leg. */
int asscpp_eval_batch_device(AsscppCtx* ctx, const double* d_poses, int64_t nviews, int covered_set,
                             uint32_t flags, const AsscppViewResult** d_results);

/* extractVisibleSurface for one pose: ascending original point indices of the visible cloud. */
int asscpp_get_visible(AsscppCtx* ctx, const double pose[7], int32_t* idx_out, int64_t cap, int64_t* n_out);

/* -- covered sets (Node::cloud_filtered / globalCloud as bitsets) ------------------------------ */
int asscpp_covered_new(AsscppCtx* ctx, int* set_out);
int asscpp_covered_free(AsscppCtx* ctx, int set);
int asscpp_covered_clear(AsscppCtx* ctx, int set);
int asscpp_covered_clone(AsscppCtx* ctx, int src, int* set_out);
int asscpp_covered_or_set(AsscppCtx* ctx, int dst, int src);
/* union of the visible sets of the given views whose visible count exceeds min_visible
 * (heuristic.cpp:344-346: "if(pts.size()>5) globalCloud += pts"; use -1 to take every view) */
int asscpp_covered_or_views(AsscppCtx* ctx, int set, const double* poses, int64_t nviews, int64_t min_visible);
int asscpp_covered_or_points(AsscppCtx* ctx, int set, const int32_t* idx, int64_t n); /* original indices */
/* coverage voxels of an arbitrary cloud (n x xyz float32, host), as calcCoveragePercent / pointCloudDiff voxelise it
 * (occlusion_culling.cpp:387-428, heuristic.cpp:92-96): OR-ed into the set's voxel bits; point bits are not touched */
int asscpp_covered_or_cloud(AsscppCtx* ctx, int set, const float* xyz, int64_t n);
/* complement of the voxel bits over the model's coverage voxels (to cluster the voxels OF a cloud instead of the
 * voxels a set lacks: clusteringPointCloud is handed the uncovered cloud itself, heuristic.cpp:147) */
int asscpp_covered_invert_voxels(AsscppCtx* ctx, int set);
int asscpp_covered_popcount(AsscppCtx* ctx, int set, int64_t* n_points, int64_t* n_voxels);
/* host bitsets, bit i of 64-bit word i/64: points in ORIGINAL index order; voxels by compact id
 * (ascending k,j,i order of the model's coverage grid) */
int asscpp_covered_export(AsscppCtx* ctx, int set, uint64_t* point_bits, uint64_t* voxel_bits);
int asscpp_covered_import(AsscppCtx* ctx, int set, const uint64_t* point_bits, const uint64_t* voxel_bits);
/* calcCoveragePercent (occlusion_culling.cpp:375-444) of a covered set, float32 arithmetic as :431-432 */
int asscpp_coverage_percent(AsscppCtx* ctx, int set, float* pct);
/* calcCoveragePercent of an arbitrary cloud (n x xyz float32, host memory) */
int asscpp_coverage_percent_cloud(AsscppCtx* ctx, const float* xyz, int64_t n, float* pct);

/* -- uncovered regions of the adaptive loop (coverage_path_planning_heuristic.cpp:88-246) ------------------ */
/* pointCloudDiff (:88-145) in bitset form: the model's coverage voxels NOT in the set, in the reference's k, j, i
 * scan order (= ascending compact id), with their cell centres (getCentroidCoordinate).  ids / centres_xyz may be
 * NULL; at most cap entries are written; *n_out receives the full count (uncovered % = 100 * n / n_cov_voxels). */
int asscpp_uncovered_voxels(AsscppCtx* ctx, int set, int32_t* ids, float* centres_xyz, int64_t cap, int64_t* n_out);
/* The same with the reference's quirk (flags & ASSCPP_DIFF_PCL_ALIASING): pointCloudDiff looks every model cell up in
 * the COVERED cloud's voxel grid through PCL's getCentroidIndexAt (heuristic.cpp:120-123), which linearises the cell
 * with that grid's own box and only checks 0 <= idx < size; a model cell outside the box whose index is that of a
 * covered cell is therefore not reported.  With the flag the list is exactly the reference's. */
enum { ASSCPP_DIFF_PCL_ALIASING = 1u };
int asscpp_uncovered_voxels_ex(AsscppCtx* ctx, int set, uint32_t flags, int32_t* ids, float* centres_xyz, int64_t cap,
                               int64_t* n_out);

/* Allocation-free variant of the uncovered-id extraction.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  The count -> scan -> emit kernels of asscpp_uncovered_voxels run with their per-word
 * offsets in a static device array of ASSCPP_NOALLOC_MAX_VOXELS / 32 + 1 int32 (1 MiB, held by every GPU that loads
 * the library until exit) instead of a context buffer; the call needs no context, allocates nothing (no cudaMalloc,
 * cudaFree or allocating library routine) and does not synchronise.
 *   d_voxel_bits  DEVICE pointer, (n_voxels + 31) / 32 words: bit i set = voxel i covered (the voxel bitset of
 *                 asscpp_covered_export, as 32-bit words)
 *   d_ids         DEVICE pointer, cap entries: the ids whose bit is clear, ascending (at most cap are written)
 *   d_count       DEVICE pointer to one int64: the full count of clear bits
 *   stream        cudaStream_t of the CURRENT device the three launches go on (NULL = default stream)
 *   err, err_len  optional host buffer that receives the error text
 * n_voxels > ASSCPP_NOALLOC_MAX_VOXELS is refused with ASSCPP_ERANGE.  All calls on a device share the one array:
 * calls that overlap on the GPU (different streams, or threads launching concurrently) race; calls on one stream,
 * or otherwise ordered, do not.  Nothing in the array has to start at zero: each call rewrites every word it reads. */
#define ASSCPP_NOALLOC_MAX_VOXELS (1 << 23)
int asscpp_uncovered_ids_noalloc(const uint32_t* d_voxel_bits, int64_t n_voxels, int32_t* d_ids, int64_t cap,
                                 int64_t* d_count, void* stream, char* err, int64_t err_len);

/* clusteringPointCloud (:147-190, pcl::EuclideanClusterExtraction: tolerance 0.8, sizes 5..50000 in the reference)
 * over those voxel centres + findClusterBB (:218-246) per cluster.  Clusters come largest first (equal sizes: by
 * smallest member id).  labels: n_cov_voxels entries, cluster index or -1 (covered, or cluster size out of range);
 * may be NULL. */
typedef struct AsscppCluster {
  int32_t size;        /* voxels in the cluster                                   */
  int32_t first_id;    /* smallest compact coverage-voxel id of the cluster       */
  int32_t min_ijk[3];  /* lattice box of the cluster                              */
  int32_t max_ijk[3];
  double grid_start[3]; /* findClusterBB gridStart.position (:242-244)            */
  double grid_size[3];  /* findClusterBB gridSize (:238-240)                      */
} AsscppCluster;
int asscpp_uncovered_clusters(AsscppCtx* ctx, int set, double tolerance, int32_t min_size, int32_t max_size,
                              int32_t* labels, AsscppCluster* clusters, int64_t cap, int64_t* n_clusters);

/* -- geometric sample filters against the structure's triangle mesh (coverage_path_planning_heuristic.cpp:300-335;
 *    the reference uses a CGAL AABB tree over Simple_cartesian<double> triangles) ------------------------------- */
typedef struct AsscppMesh AsscppMesh;
/* vertices: nv x {x y z} float32 as parsed by loadOBJFile (:1355-1451); triangles: nt x 3 vertex indices (0-based).
 * Collinear triangles are dropped as the loader does (:1409, :1430). */
int asscpp_mesh_create(int device, const float* vertices, int64_t nv, const int32_t* triangles, int64_t nt,
                       AsscppMesh** out);
int asscpp_mesh_destroy(AsscppMesh* mesh);
int64_t asscpp_mesh_num_triangles(const AsscppMesh* mesh);
int64_t asscpp_mesh_launch_count(const AsscppMesh* mesh);
const char* asscpp_mesh_last_error(const AsscppMesh* mesh);
/* isFilteringConditionSatisfied's geometric part (:319-333) for n candidate positions (n x 3 doubles, host):
 * n_hits = triangles crossed by the ray from the position through ray_through (the reference uses (100, 10, 56)),
 * sq_dist = squared distance to the mesh, pass = n_hits even && min_dist^2 <= sq_dist <= max_dist^2.
 * Any output may be NULL. */
int asscpp_mesh_filter(AsscppMesh* mesh, const double* positions, int64_t n, const double ray_through[3],
                       double min_dist, double max_dist, int32_t* n_hits, double* sq_dist, uint8_t* pass);
/* isConnectionConditionSatisfied (:300-315): free_out[i] = the segment a[i] -> b[i] crosses no triangle */
int asscpp_mesh_segments_free(AsscppMesh* mesh, const double* a, const double* b, int64_t n, uint8_t* free_out);

/* -- streams / timing -------------------------------------------------------------------------- */
int asscpp_sync(AsscppCtx* ctx);
void* asscpp_get_stream(AsscppCtx* ctx); /* cudaStream_t all kernels are launched on */
int asscpp_timer_begin(AsscppCtx* ctx);           /* cudaEventRecord on that stream */
int asscpp_timer_end(AsscppCtx* ctx, float* ms);  /* record + synchronise + elapsed  */
/* kernels launched by this context since creation (the bench's gpu_launches claim) */
int64_t asscpp_launch_count(const AsscppCtx* ctx);
/* developer aid: SM-cycle counters of eval_views_kernel's phases, summed over views, read and reset.
 * out[0..3] = cull, grid build, DDA, point visit; out[4] = views; out[5] = CTA lifetimes; out[6..15] = finer
 * splits of the cull (see tools/phase_profile.py).  enable != 0
 * turns the counters on for subsequent evaluations, 0 turns them off. */
int asscpp_debug_phase_cycles(AsscppCtx* ctx, int enable, uint64_t out[16]);
/* device time of the dominant kernel (eval_views) accumulated between timer_begin/timer_end */
int asscpp_eval_kernel_ms(AsscppCtx* ctx, float* ms, int64_t* launches);

/* -- multi-GPU: one context per rank, NCCL over NVLink ---------------------------------------- */
int asscpp_comm_unique_id(uint8_t id_out[128]);
int asscpp_comm_init(AsscppCtx* ctx, int nranks, int rank, const uint8_t id[128]);
/* gain-score allgather: every rank contributes its nviews_local results; out (host) receives
 * nranks * nviews_local results in rank order.  out == NULL keeps the gathered scores on the
 * device and returns without synchronising. */
int asscpp_allgather_results(AsscppCtx* ctx, int64_t nviews_local, AsscppViewResult* out);
/* The same for the 8 bytes a search loop needs per view: gain_pts / gain_vox (host, nranks * nviews_local entries
 * each, rank order; either may be NULL).  Both NULL: the gathered scores stay on the device, no synchronisation.
 * EVERY RANK MUST PASS THE SAME nviews_local (pad uneven shards with repeated poses): the host-output forms of both
 * allgathers exchange the counts first and return ASSCPP_EINVAL on a mismatch; the device-output forms do not check. */
int asscpp_allgather_gain(AsscppCtx* ctx, int64_t nviews_local, uint32_t* gain_pts, uint32_t* gain_vox);
/* covered-bitset OR-allreduce (NCCL has no OR op: allgather + fused OR kernel) */
int asscpp_allreduce_or(AsscppCtx* ctx, int set);
/* device time spent in the two collectives between asscpp_timer_begin / asscpp_timer_end (CUDA events around them) */
int asscpp_collective_ms(AsscppCtx* ctx, float* allgather_ms, int64_t* allgather_calls, float* allreduce_or_ms,
                         int64_t* allreduce_or_calls);
int asscpp_comm_destroy(AsscppCtx* ctx);
/* one process, several GPUs: a communicator over n contexts of THIS process, one per device (ncclCommInitAll);
 * context i becomes rank i.  The collectives above are then called once per context, from one host thread each. */
int asscpp_comm_init_all(AsscppCtx** ctxs, int n);

/* -- single-process multi-GPU evaluator -------------------------------------------------------------------------
 * The reference is one process (coverage_heuristic_test.cpp:42): a program that links the C++ shim cannot start a
 * rank per GPU.  AsscppMulti holds one context per listed device (model replicated), one host worker thread per
 * device, and shards each batch over them in contiguous blocks of whole groups; covered sets are mirrored (the same
 * handle on every device) and kept equal by the NCCL OR-allreduce. */
typedef struct AsscppMulti AsscppMulti;
int asscpp_multi_create(const int* devices, int ndev, const float* xyz, int64_t n, const AsscppParams* params,
                        AsscppMulti** out);
int asscpp_multi_destroy(AsscppMulti* m);
int asscpp_multi_num_devices(const AsscppMulti* m);
AsscppCtx* asscpp_multi_context(AsscppMulti* m, int i);
const char* asscpp_multi_last_error(const AsscppMulti* m);
int asscpp_multi_covered_new(AsscppMulti* m, int* set_out);
int asscpp_multi_covered_free(AsscppMulti* m, int set);
int asscpp_multi_covered_clear(AsscppMulti* m, int set);
int asscpp_multi_covered_popcount(AsscppMulti* m, int set, int64_t* n_points, int64_t* n_voxels);
int asscpp_multi_coverage_percent(AsscppMulti* m, int set, float* pct);
int asscpp_multi_allreduce_or(AsscppMulti* m, int set);
/* asscpp_eval_ex over all devices (host poses only).  results / gain_pts / gain_vox (any may be NULL): nviews entries
 * in the order of poses.  opts->accum_set >= 0: the accumulator is OR-allreduced afterwards. */
int asscpp_multi_eval(AsscppMulti* m, const double* poses, int64_t nviews, const AsscppEvalOpts* opts,
                      AsscppViewResult* results, uint32_t* gain_pts, uint32_t* gain_vox);
int asscpp_multi_covered_or_views(AsscppMulti* m, int set, const double* poses, int64_t nviews, int64_t min_visible);

/* -- the adaptive sampling loop, host side in C++ (sspp/src/pathplanner.cpp:106-201, 729-850) ---------------------- */
/* PathPlanner::generateRegularGrid's candidate lattice: robot poses {x y z qx qy qz qw} in its loop order (z, x, y,
 * then 360 / orientation_res_deg yaws from 0).  Returns the count (robot_poses may be NULL / cap 0 to ask for it). */
int64_t asscpp_grid_candidates(const double grid_start[3], const double grid_size[3], double res,
                               double orientation_res_deg, double* robot_poses, int64_t cap);
typedef struct AsscppSamplingOpts {
  double grid_start[3], grid_size[3];   /* first stage (coverage_heuristic_test.cpp:75-80)                           */
  double start_res, res_decrement;      /* 4.5, 1.5 in the demo (:137)                                                */
  double orientation_res_deg;           /* 45 (pathplanner.cpp:736)                                                   */
  double min_dist, max_dist;            /* stand-off window of the mesh filter: 1, 4 (pathplanner.cpp:160)            */
  double stop_uncovered_pct;            /* 0.01 (pathplanner.cpp:776)                                                 */
  int32_t n_sensors, reserved;          /* <= 4                                                                       */
  double sensor_offset[4][3];           /* mounts (coverage_heuristic_test.cpp:94-95), as float32 values              */
  double sensor_rpy[4][3];
} AsscppSamplingOpts;
typedef struct AsscppSamplingStage {
  double res, uncovered_pct_before, uncovered_pct_after;
  int64_t grids, candidates, passed_mesh, accepted, clusters;
} AsscppSamplingStage;
/* PathPlanner::dynamicNodesGenerationAndConnection without the node connection: per stage ONE mesh launch and ONE
 * evaluation launch over all GPUs of m for every candidate of every cluster grid, the accepted sensors' clouds OR-ed
 * into global_set (a set of m), then pointCloudDiff / clusteringPointCloud / findClusterBB on the device.
 * robot_poses (cap x 7) / sensor_poses (cap x n_sensors x 7) receive the accepted waypoints in the search space's
 * order (newest first, duplicates dropped as SearchSpace::insertNode does); either may be NULL. */
int asscpp_adaptive_sampling(AsscppMulti* m, AsscppMesh* mesh, const AsscppSamplingOpts* opts, int global_set,
                             double* robot_poses, double* sensor_poses, int64_t cap, int64_t* n_out,
                             AsscppSamplingStage* stages, int stage_cap, int* n_stages);

/* -- synthetic workloads of BASELINE.json configs 2/3/5 (host generators) ------- */
/* These host-only helpers (generators, pose transforms) are also exported by asscpp_b200/lib/libasscpp_host.so,
 * which has no CUDA in it: the benchmark's CPU reference arm draws its inputs from there, so it never maps the
 * product's CUDA library. */
/* structure_seed places the shapes (default 0xA55C0001); point_seed / view_seed drive the
 * sampling streams (defaults 0xA55C0002 / 0xA55C0003).  Deterministic, host only. */
int asscpp_synth_structure_v1(uint64_t structure_seed, uint64_t point_seed, int64_t n, float* xyz_out);
int asscpp_synth_views_v1(uint64_t structure_seed, uint64_t view_seed, int64_t nviews, double* poses_out);

/* sspp/src/sensors.cpp:47-90 robot pose -> sensor pose (batcher input), host, double */
int asscpp_robot2sensor(const double robot_pose[7], const double offset[3], const double rpy[3],
                        double sensor_pose[7]);
/* the same for n robot poses at once (robot_poses, sensor_poses: n x 7 doubles): the batcher's input transform for a
 * whole search space (pathplanner.cpp:122-147 calls it once per waypoint and sensor) */
int asscpp_robot2sensor_batch(const double* robot_poses, int64_t n, const double offset[3], const double rpy[3],
                              double* sensor_poses);
/* tf::Matrix3x3(q).getRPY(roll, pitch, yaw) of a pose, as used by occlusion_culling.cpp:497-500, 558-561 (host) */
int asscpp_pose_rpy(const double pose[7], double rpy[3]);
/* OcclusionCulling::pointCloudViewportTransform (occlusion_culling.cpp:538-618), host: n x xyz float32 -> same */
int asscpp_viewport_transform(const double pose[7], const float* xyz, int64_t n, float* xyz_out);
/* frustum planes (l r t b f n) and camera origin of a pose: frustum_culling.cpp:75-141 */
int asscpp_frustum_planes(const AsscppParams* p, const double pose[7], float planes[24], float origin[3]);

#ifdef __cplusplus
}
#endif
#endif
