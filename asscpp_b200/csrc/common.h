// common.h -- structures shared by the host side (abi.cu, preprocess.cpp) and the kernels.
#pragma once
#include <stdint.h>
#include <vector_types.h>

#include "../../include/asscpp_gpu.h"

namespace asscpp {

#ifndef ASSCPP_CHUNK_MAX_POINTS
#define ASSCPP_CHUNK_MAX_POINTS 128
#endif
constexpr int kChunkMaxPoints = ASSCPP_CHUNK_MAX_POINTS;  // leaf bucket of the cull hierarchy (whole voxels)
constexpr int kSuperFanout = 32;      // chunks per super-chunk (one warp ballot)

// Frustum constants that depend only on the parameters; evaluated once on the host with libm so
// the double-precision tan() is the same one the CPU path uses (frustum_culling.cpp:88-95).
struct FrustumConst {
  float np_h, np_w, fp_h, fp_w;
  float near_dist, far_dist;
};

// Device-resident model, laid out once by asscpp_create.
struct ModelDev {
  int64_t n;            // points
  int64_t n_vox;        // occupied voxels (occlusion lattice)
  int64_t n_cov;        // occupied voxels (coverage lattice)
  int64_t n_chunks;
  int64_t n_super;
  const float4* pts;    // [n] x y z + voxel id bits, sorted by (morton(voxel), original index)
  const int32_t* orig;  // [n] original index of sorted position
  const int32_t* pos_of;  // [n] sorted position of original index
  const int4* vox;      // [n_vox+1] {i, j, k, first point}; sentinel holds n
  const int32_t* vox_cov;  // [n_vox] coverage id shared by all points of the voxel, -1 if they straddle coverage cells
  const int32_t* pt_cov;   // [n] coverage id per sorted position
  const float4* chunk_lo;  // [n_chunks] aabb min + first point (bits)
  const float4* chunk_hi;  // [n_chunks] aabb max + end point (bits)
  const int32_t* chunk_vox;  // [n_chunks+1] first voxel of the chunk
  const float4* super_lo;  // [n_super]
  const float4* super_hi;  // [n_super]
  float leaf, inv_leaf, half_leaf;
  float pcd_origin[3];
  float max_abs_coord;  // bound on |x|, |y|, |z| over the model
  float dist_thr_sq;    // largest f32 s with sqrt((double)s) <= (double)(leaf*2): occluder iff sum > this
  int cov_same;         // coverage lattice == occlusion lattice
  int32_t min_b[3], max_b[3];
};

struct ViewDev {  // per view, written by planes_kernel
  float planes[24];  // l r t b f n
  float origin[3];
  float pad;
};

// pointCloudViewportTransform constants of one view (occlusion_culling.cpp:545-584)
struct ViewTrig {
  double t[3];            // -camera position
  double cz, sz, cy, sy;  // cos / sin of -yaw and of -pitch
  double pad;
};

// accuracy accumulators of one view: fixed-point sums (value * 2^S split into hi * 2^32 + lo) and ordered-int keys
struct AccRaw {
  unsigned long long d_hi, d_lo, z_hi, z_lo;
  int32_t min_key, max_key;
  uint32_t visible, pad;
};

struct CoveredDev {
  uint32_t* pts;  // bit per sorted position
  uint32_t* vox;  // bit per compact coverage id
};

constexpr int kSmallMaxViews = 512;   // most views eval_views_small_kernel takes (poses are kernel parameters)
constexpr int kSmallMaxGroup = 8;     // largest group whose union gains it computes in the kernel

struct EvalArgs {
  ModelDev m;
  const ViewDev* views;
  int64_t nviews;
  int32_t* work_counter;     // device scalar, zeroed before launch
  AsscppViewResult* results;
  uint2* gains;              // [nviews] {gain_pts, gain_vox}: the 8 bytes per view a search loop reads back
  int32_t* error_count;      // views whose scratch bound was violated (kVsError), zeroed before launch
  CoveredDev covered;        // score against (pts == nullptr: empty set)
  CoveredDev accum;          // OR visible sets into (pts == nullptr: none)
  int64_t accum_min_visible;
  uint32_t flags;
  const ViewTrig* trig;      // per view, ASSCPP_EVAL_ACCURACY only
  AccRaw* acc;               // per view, ASSCPP_EVAL_ACCURACY only
  double acc_scale_d, acc_scale_z;  // 2^S of the two fixed-point sums
  // per-CTA scratch (global memory, L2-resident in steady state)
  uint2* chunk_list;         // [ctas][chunk_cap] candidate chunks {first point, end point}
  int64_t chunk_cap;
  uint32_t* vox_list;        // [ctas][list_cap] voxels occupied for the view (0xffffffff: duplicate)
  int64_t list_cap;
  uint32_t* vis_list;        // [ctas][vis_cap] voxels whose ray came back free (second gate applied in P4b)
  int64_t vis_cap;
  uint4* ray_buf;            // [ctas][3][ray_cap] prepared ray records (SoA of three 16-byte planes)
  int64_t ray_cap;
  int32_t walk_thresh;       // leave the DDA burst loop for service/refill when fewer lanes than this still walk
  uint32_t* grid_scratch;    // [ctas][grid_cap_words] bit-grid when it does not fit shared memory
  int64_t grid_cap_words;
  uint32_t* cov_scratch;     // [ctas][cov_words] per-view coverage-voxel marks
  int64_t cov_words;
  int halo_iters;            // ... plus at most this many more while the ray is still in the halo layer
  int setup_iters;           // DDA iterations done in the converged ray setup before a ray is queued
  int group;                 // views per candidate (sensors of one node): cov_scratch is kept across a group
  int cov_direct;            // coverage lattice == occlusion lattice and group == 1: no marks needed
  uint32_t* view_cov;        // split-group mode (small grouped batches): [nviews][cov_words] coverage-voxel bitset of
                             // every view's visible cloud; every view is its own work unit and group_gain_kernel
                             // turns the bitsets into the per-view union gains afterwards.  nullptr: off
  int split_group;           // views per group in split-group mode (A.group is 1 then)
  uint8_t* vstate;           // [ordered ctas][n_vox] per-voxel state for the ordered (Q5) path
  int32_t* ordered_list;     // views that saw a Q5 carry
  int32_t* ordered_count;
  uint32_t smem_grid_words;  // capacity of the shared-memory grid
  // small batches (the A* use), one launch for the whole evaluation: poses arrive as kernel parameters and the planes are
  // computed by the CTA that takes the view; the last view of a group to finish computes the group's union gains; gains
  // and counters are mirrored into mapped host memory; the last CTA resets the device counters for the next launch
  int chunk_cull;            // coarse lattice: boundary chunks are culled chunk by chunk (no per-voxel pass)
  int small;                 // 1: eval_views_small_kernel
  int32_t epoch;             // written to host_counters[2] when the launch is complete
  FrustumConst fc;
  uint2* host_gains;         // [nviews] mapped host mirror of gains
  AsscppViewResult* host_results;  // [nviews] ... of results
  AccRaw* host_acc;          // [nviews] ... of acc (ASSCPP_EVAL_ACCURACY)
  int32_t* host_counters;    // mapped: {views that need the ordered replay, views with a scratch error}
  uint32_t* group_ticket;    // [groups] views of the group that have finished (wraps to 0 by itself)
  unsigned long long* phase_cycles;  // optional profiling counters (nullptr: off)
};

constexpr uint32_t kVoxInside = 0x80000000u;  // vox_list flag: every point of the voxel passes the frustum test
constexpr uint32_t kVoxMask = 0x7fffffffu;
constexpr uint32_t kVsError = 0x80000000u;  // internal consistency failure (scratch overflow)

}  // namespace asscpp
