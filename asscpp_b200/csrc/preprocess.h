// preprocess.h -- one-time host layout of the model cloud (replaces the constructor work at
// cscpp/src/occlusion_culling.cpp:26-79: voxelFilterOriginal + OriginalVoxelsSize).
#pragma once
#include <string>
#include <vector>

#include "common.h"

namespace asscpp {

struct Float4 { float x, y, z, w; };
struct Int4 { int32_t x, y, z, w; };

struct ModelHost {
  int64_t n = 0, n_vox = 0, n_cov = 0, n_chunks = 0, n_super = 0;
  std::vector<Float4> pts;
  std::vector<int32_t> orig, pos_of, vox_cov, pt_cov, chunk_vox;
  std::vector<Int4> vox;
  std::vector<Float4> chunk_lo, chunk_hi, super_lo, super_hi;
  std::vector<int64_t> cov_keys;  // sorted linear keys of the coverage voxels (compact id = rank)
  int32_t min_b[3] = {0, 0, 0}, max_b[3] = {0, 0, 0};          // occlusion lattice
  int32_t cov_min_b[3] = {0, 0, 0}, cov_div[3] = {1, 1, 1};    // coverage lattice
  float inv_leaf = 0, inv_cov = 0, dist_thr_sq = 0;
  int cov_same = 0;
};

int preprocess_model(const float* xyz, int64_t n, const AsscppParams& p, ModelHost* out, std::string* err);

// frustum helpers (host; same arithmetic as planes_kernel)
FrustumConst frustum_const(const AsscppParams& p);
void frustum_planes_host_fc(const FrustumConst& fc, const double pose[7], float planes[24], float origin[3],
                            float corners[24]);
void frustum_planes_host(const AsscppParams& p, const double pose[7], float planes[24], float origin[3],
                         float corners[24]);
// pointCloudViewportTransform constants (occlusion_culling.cpp:545-584) with the host's libm
void viewport_trig_host(const double pose[7], ViewTrig* out);
void pose_rpy_host(const double pose[7], double rpy[3]);
// largest extent of the frustum along any axis over all orientations (bound for grid scratch)
double frustum_max_extent(const AsscppParams& p);

}  // namespace asscpp
