// kernels.h -- host-callable launchers of kernels.cu
#pragma once
#include <cuda_runtime.h>

#include "common.h"

namespace asscpp {

size_t eval_smem_bytes(uint32_t grid_words);
size_t eval_smem_header_bytes();
int ray_queue_cap();
cudaError_t configure_kernels(size_t smem_bytes);

void launch_planes(const double* d_poses, int64_t nviews, const FrustumConst& fc, ViewDev* views, cudaStream_t s);
void launch_trig(const double* d_poses, int64_t nviews, ViewTrig* out, cudaStream_t s);
void launch_eval(const EvalArgs& a, int ctas, int threads, size_t smem_bytes, cudaStream_t s);
void launch_eval_small(const EvalArgs& a, const double* h_poses, const ViewDev* h_views, int ctas, int threads,
                       size_t smem_bytes, cudaStream_t s);
int tiny_max_views();
void launch_group_gain(const EvalArgs& a, int group, cudaStream_t s);
void launch_ordered(const EvalArgs& a, int ctas, size_t smem_bytes, cudaStream_t s);
void launch_or_words(uint32_t* dst, const uint32_t* src, int64_t words, cudaStream_t s);
void launch_or_reduce(uint32_t* dst, const uint32_t* src, int64_t words, int nsrc, cudaStream_t s);
void launch_popcount(const uint32_t* bits, int64_t words, unsigned long long* out, cudaStream_t s);
void launch_permute_bits(const uint32_t* src, uint32_t* dst, const int32_t* map, int64_t n, cudaStream_t s);
void launch_or_points(CoveredDev set, const int32_t* idx, int64_t n, const ModelDev& m, cudaStream_t s);
void launch_cloud_cover(const float* xyz, int64_t n, float inv_cov, const int32_t cmin[3], const int32_t cdiv[3],
                        const int64_t* keys, int64_t nkeys, uint32_t* marks, cudaStream_t s);
// regions.cu
void launch_uncovered(const uint32_t* vox_bits, int64_t n_cov, int32_t* wcount, int64_t* d_total, const int64_t* keys,
                      const int32_t cmin[3], const int32_t cdiv[3], float leaf, int32_t* ids, float* centres,
                      int64_t cap, cudaStream_t s);
void launch_covered_alias(const uint32_t* vox_bits, uint32_t* out_bits, int64_t n_cov, const int64_t* keys,
                          const int32_t cdiv[3], int32_t* box6, cudaStream_t s);
void launch_clusters(const uint32_t* vox_bits, int64_t n_cov, const int64_t* keys, const int32_t cmin[3],
                     const int32_t cdiv[3], const int3* offs, int noffs, int32_t* parent, int32_t* root, int32_t* size,
                     int32_t* box, cudaStream_t s);
void launch_invert_bits(uint32_t* bits, int64_t nbits, cudaStream_t s);
void launch_fill_bytes(uint8_t* p, int64_t n, uint8_t v, cudaStream_t s);

}  // namespace asscpp
