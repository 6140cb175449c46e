// regions.cu -- uncovered-region extraction of the ASSCPP adaptive loop on the covered-voxel bitset
// (SURVEY.md 8f rank 2):
//   pointCloudDiff        cscpp/src/coverage_path_planning_heuristic.cpp:88-145   -> uncovered_* kernels
//   clusteringPointCloud  ...:147-190 (<PCL EuclideanClusterExtraction>)          -> cluster_* kernels
//   findClusterBB         ...:218-246                                             -> per-cluster lattice boxes
// The coverage voxels of the model are the sorted linear keys `keys` (compact id = rank, ascending k, j, i: the
// reference's scan order); a covered set is one bit per compact id.
#include <cuda_runtime.h>

#include <atomic>
#include <climits>
#include <cstdio>

#include "../../include/asscpp_gpu.h"

#include "common.h"
#include "kernels.h"

namespace asscpp {

namespace {

__device__ __forceinline__ bool vox_uncovered(const uint32_t* bits, int64_t id) {
  return !((bits[id >> 5] >> (id & 31)) & 1u);
}

// zero bits per 32-bit word of the covered set (ids >= n_cov masked off)
__global__ void uncovered_count_kernel(const uint32_t* __restrict__ bits, int64_t n_cov, int32_t* __restrict__ wcount) {
  const int64_t words = (n_cov + 31) / 32;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (int64_t)gridDim.x * blockDim.x) {
    uint32_t z = ~bits[w];
    if ((w + 1) * 32 > n_cov) z &= (n_cov & 31) ? ((1u << (n_cov & 31)) - 1u) : 0xffffffffu;
    wcount[w] = __popc(z);
  }
}

// exclusive scan of wcount in place (one CTA; the array has n_cov / 32 entries), total to *total
__global__ void uncovered_scan_kernel(int32_t* __restrict__ wcount, int64_t words, int64_t* __restrict__ total) {
  __shared__ int32_t warp_sums[32];
  __shared__ int64_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int64_t base = 0; base < words; base += blockDim.x) {
    const int64_t i = base + threadIdx.x;
    const int32_t v = i < words ? wcount[i] : 0;
    int32_t incl = v;
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int32_t ws = lane < nwarps ? warp_sums[lane] : 0, wi = ws;
      for (int o = 1; o < 32; o <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= o) wi += t;
      }
      warp_sums[lane] = wi - ws;  // exclusive
    }
    __syncthreads();
    const int64_t c = carry;
    if (i < words) wcount[i] = (int32_t)(c + warp_sums[warp] + incl - v);
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = c + warp_sums[warp] + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}

// heuristic.cpp:108-140: ids in ascending order with getCentroidCoordinate of each cell
// (voxel_grid_occlusion_estimation.h:143-157: (b_min + leaf / 2) + float(ijk - min_b) * leaf, float32, unfused)
__global__ void uncovered_emit_kernel(const uint32_t* __restrict__ bits, int64_t n_cov, const int32_t* __restrict__ woff,
                                      const long long* __restrict__ keys, int3 cmin, int3 cdiv, float leaf,
                                      int32_t* __restrict__ ids, float* __restrict__ centres, int64_t cap) {
  const int64_t words = (n_cov + 31) / 32;
  for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (int64_t)gridDim.x * blockDim.x) {
    uint32_t z = ~bits[w];
    if ((w + 1) * 32 > n_cov) z &= (n_cov & 31) ? ((1u << (n_cov & 31)) - 1u) : 0xffffffffu;
    int64_t at = woff[w];
    while (z) {
      const int b = __ffs(z) - 1;
      z &= z - 1;
      const int64_t id = w * 32 + b;
      if (at < cap) {
        if (ids) ids[at] = (int32_t)id;
        if (centres) {
          const long long key = keys[id];
          const int r[3] = {(int)(key % cdiv.x), (int)((key / cdiv.x) % cdiv.y), (int)(key / ((long long)cdiv.x * cdiv.y))};
          const int mn[3] = {cmin.x, cmin.y, cmin.z};
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            const float b_min = __fmul_rn((float)mn[k], leaf);
            centres[3 * at + k] = __fadd_rn(__fadd_rn(b_min, __fmul_rn(leaf, 0.5f)), __fmul_rn((float)r[k], leaf));
          }
        }
      }
      ++at;
    }
  }
}

// ---- quirk Q11: pointCloudDiff's unchecked linear index ----------------------------------------------------------
// heuristic.cpp:120-123 looks every model cell up in the COVERED cloud's voxel grid with <PCL getCentroidIndexAt>, which
// linearises (ijk - min_b) . (1, div0, div0 div1) with the covered grid's own box and checks only 0 <= idx < size.  A
// model cell outside that box can land on the index of another, covered cell and is then NOT reported as uncovered.
// box[0..2] = min, box[3..5] = max of the covered cells (lattice coordinates relative to the model grid).
__global__ void covered_box_kernel(const uint32_t* __restrict__ bits, int64_t n_cov, const long long* __restrict__ keys,
                                   int3 cdiv, int32_t* __restrict__ box) {
  int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_cov; i += (int64_t)gridDim.x * blockDim.x) {
    if (vox_uncovered(bits, i)) continue;
    const long long key = keys[i];
    const int r[3] = {(int)(key % cdiv.x), (int)((key / cdiv.x) % cdiv.y), (int)(key / ((long long)cdiv.x * cdiv.y))};
#pragma unroll
    for (int k = 0; k < 3; ++k) { mn[k] = min(mn[k], r[k]); mx[k] = max(mx[k], r[k]); }
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int lo = __reduce_min_sync(0xffffffffu, mn[k]), hi = __reduce_max_sync(0xffffffffu, mx[k]);
    if ((threadIdx.x & 31) == 0 && lo <= hi) { atomicMin(&box[k], lo); atomicMax(&box[3 + k], hi); }
  }
}

// out = bits, plus every uncovered model cell whose linear index in the covered grid is that of a covered cell
__global__ void covered_alias_kernel(const uint32_t* __restrict__ bits, uint32_t* __restrict__ out, int64_t n_cov,
                                     const long long* __restrict__ keys, int3 cdiv, const int32_t* __restrict__ box) {
  const int mn0 = box[0], mn1 = box[1], mn2 = box[2];
  if (mn0 > box[3]) return;  // empty covered set: an empty leaf layout answers -1 to every lookup
  const int d0 = box[3] - mn0 + 1, d1 = box[4] - mn1 + 1, d2 = box[5] - mn2 + 1;
  const int size = d0 * d1 * d2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_cov; i += (int64_t)gridDim.x * blockDim.x) {
    if (!vox_uncovered(bits, i)) continue;
    const long long key = keys[i];
    const int r0 = (int)(key % cdiv.x) - mn0, r1 = (int)((key / cdiv.x) % cdiv.y) - mn1,
              r2 = (int)(key / ((long long)cdiv.x * cdiv.y)) - mn2;
    const int lin = r0 + r1 * d0 + r2 * d0 * d1;  // <PCL> (ijk - min_b_).dot(divb_mul_), int arithmetic
    if (lin < 0 || lin >= size) continue;
    const int a2 = lin / (d0 * d1), rem = lin - a2 * d0 * d1, a1 = rem / d0, a0 = rem - a1 * d0;
    const long long akey = (a0 + mn0) + (long long)cdiv.x * ((a1 + mn1) + (long long)cdiv.y * (a2 + mn2));
    int64_t lo = 0, hi = n_cov;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (keys[mid] < akey) lo = mid + 1; else hi = mid;
    }
    if (lo < n_cov && keys[lo] == akey && !vox_uncovered(bits, lo)) atomicOr(&out[i >> 5], 1u << (i & 31));
  }
}

// ---- clusters: lock-free union-find over the uncovered voxels, edges = lattice offsets within the tolerance ----
__global__ void cluster_init_kernel(const uint32_t* __restrict__ bits, int64_t n_cov, int32_t* __restrict__ parent) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_cov; i += (int64_t)gridDim.x * blockDim.x)
    parent[i] = vox_uncovered(bits, i) ? (int32_t)i : -1;
}

__device__ __forceinline__ int32_t uf_find(const int32_t* parent, int32_t x) {
  int32_t p = ((volatile const int32_t*)parent)[x];
  while (p != x) { x = p; p = ((volatile const int32_t*)parent)[x]; }
  return x;
}

__global__ void cluster_union_kernel(const uint32_t* __restrict__ bits, int64_t n_cov, const long long* __restrict__ keys,
                                     int3 cdiv, const int3* __restrict__ offs, int noffs, int32_t* parent) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_cov; i += (int64_t)gridDim.x * blockDim.x) {
    if (!vox_uncovered(bits, i)) continue;
    const long long key = keys[i];
    const int a = (int)(key % cdiv.x), b = (int)((key / cdiv.x) % cdiv.y), c = (int)(key / ((long long)cdiv.x * cdiv.y));
    for (int o = 0; o < noffs; ++o) {
      const int3 d = offs[o];
      const int na = a + d.x, nb = b + d.y, nc = c + d.z;
      if (na < 0 || nb < 0 || nc < 0 || na >= cdiv.x || nb >= cdiv.y || nc >= cdiv.z) continue;
      const long long nkey = na + (long long)cdiv.x * (nb + (long long)cdiv.y * nc);
      if (nkey > key) continue;  // every undirected edge once, from its larger end
      int64_t lo = 0, hi = n_cov;
      while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (keys[mid] < nkey) lo = mid + 1; else hi = mid;
      }
      if (lo >= n_cov || keys[lo] != nkey || !vox_uncovered(bits, lo)) continue;
      // hook the larger root under the smaller one; retry when another thread moved it first
      int32_t x = (int32_t)i, y = (int32_t)lo;
      for (;;) {
        x = uf_find(parent, x);
        y = uf_find(parent, y);
        if (x == y) break;
        if (x < y) { const int32_t t = x; x = y; y = t; }
        const int32_t old = atomicMin(&parent[x], y);
        if (old == x) break;
        x = old;
      }
    }
  }
}

// root of every uncovered voxel, cluster sizes and lattice boxes per root
__global__ void cluster_label_kernel(int64_t n_cov, const long long* __restrict__ keys, int3 cmin, int3 cdiv,
                                     const int32_t* __restrict__ parent, int32_t* __restrict__ root,
                                     int32_t* __restrict__ size, int32_t* __restrict__ box) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_cov; i += (int64_t)gridDim.x * blockDim.x) {
    if (parent[i] < 0) { root[i] = -1; continue; }
    const int32_t r = uf_find(parent, (int32_t)i);
    root[i] = r;
    atomicAdd(&size[r], 1);
    const long long key = keys[i];
    const int a = (int)(key % cdiv.x) + cmin.x, b = (int)((key / cdiv.x) % cdiv.y) + cmin.y,
              c = (int)(key / ((long long)cdiv.x * cdiv.y)) + cmin.z;
    atomicMin(&box[6 * (int64_t)r + 0], a); atomicMin(&box[6 * (int64_t)r + 1], b); atomicMin(&box[6 * (int64_t)r + 2], c);
    atomicMax(&box[6 * (int64_t)r + 3], a); atomicMax(&box[6 * (int64_t)r + 4], b); atomicMax(&box[6 * (int64_t)r + 5], c);
  }
}

__global__ void cluster_box_init_kernel(int64_t n_cov, int32_t* __restrict__ size, int32_t* __restrict__ box) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_cov; i += (int64_t)gridDim.x * blockDim.x) {
    size[i] = 0;
    box[6 * i + 0] = box[6 * i + 1] = box[6 * i + 2] = INT_MAX;
    box[6 * i + 3] = box[6 * i + 4] = box[6 * i + 5] = INT_MIN;
  }
}

inline int grid_for(int64_t n) {
  int64_t b = (n + 255) / 256;
  return (int)(b < 1 ? 1 : (b > 1184 ? 1184 : b));
}

}  // namespace

void launch_uncovered(const uint32_t* vox_bits, int64_t n_cov, int32_t* wcount, int64_t* d_total, const int64_t* keys,
                      const int32_t cmin[3], const int32_t cdiv[3], float leaf, int32_t* ids, float* centres,
                      int64_t cap, cudaStream_t s) {
  if (n_cov <= 0) { cudaMemsetAsync(d_total, 0, sizeof(int64_t), s); return; }
  const int64_t words = (n_cov + 31) / 32;
  uncovered_count_kernel<<<grid_for(words), 256, 0, s>>>(vox_bits, n_cov, wcount);
  uncovered_scan_kernel<<<1, 1024, 0, s>>>(wcount, words, (int64_t*)d_total);
  uncovered_emit_kernel<<<grid_for(words), 256, 0, s>>>(vox_bits, n_cov, wcount, (const long long*)keys,
                                                        make_int3(cmin[0], cmin[1], cmin[2]),
                                                        make_int3(cdiv[0], cdiv[1], cdiv[2]), leaf, ids, centres, cap);
}

void launch_covered_alias(const uint32_t* vox_bits, uint32_t* out_bits, int64_t n_cov, const int64_t* keys,
                          const int32_t cdiv[3], int32_t* box6, cudaStream_t s) {
  if (n_cov <= 0) return;
  static const int32_t init[6] = {INT_MAX, INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MIN};
  cudaMemcpyAsync(box6, init, sizeof(init), cudaMemcpyHostToDevice, s);
  const int3 dv = make_int3(cdiv[0], cdiv[1], cdiv[2]);
  cudaMemcpyAsync(out_bits, vox_bits, (size_t)((n_cov + 31) / 32) * 4, cudaMemcpyDeviceToDevice, s);
  covered_box_kernel<<<grid_for(n_cov), 256, 0, s>>>(vox_bits, n_cov, (const long long*)keys, dv, box6);
  covered_alias_kernel<<<grid_for(n_cov), 256, 0, s>>>(vox_bits, out_bits, n_cov, (const long long*)keys, dv, box6);
}

// ---- allocation-free variant of the uncovered-id extraction (asscpp_uncovered_ids_noalloc) ------------------------
// No counterpart in the reference; see include/asscpp_gpu.h.  The per-word counts / offsets the count -> scan -> emit
// kernels need live in this file-scope array instead of a cudaMalloc'ed buffer; every GPU that loads the library holds
// it until exit (ASSCPP_NOALLOC_MAX_VOXELS / 32 + 1 words).  g_noalloc_scratch is a host-side stand-in for the symbol
// here: host code only ever uses the device address cudaGetSymbolAddress returns, fetched once per device.
__device__ __align__(256) static int32_t g_noalloc_scratch[ASSCPP_NOALLOC_MAX_VOXELS / 32 + 1];

namespace {
constexpr int kNoallocMaxDevices = 64;
std::atomic<int32_t*> g_noalloc_ptr[kNoallocMaxDevices];

void noalloc_error(char* err, int64_t err_len, const char* fmt, long long a, long long b, const char* text) {
  if (err && err_len > 0) std::snprintf(err, (size_t)err_len, fmt, a, b, text ? text : "");
}
}  // namespace

void launch_clusters(const uint32_t* vox_bits, int64_t n_cov, const int64_t* keys, const int32_t cmin[3],
                     const int32_t cdiv[3], const int3* offs, int noffs, int32_t* parent, int32_t* root, int32_t* size,
                     int32_t* box, cudaStream_t s) {
  if (n_cov <= 0) return;
  const int g = grid_for(n_cov);
  const int3 dv = make_int3(cdiv[0], cdiv[1], cdiv[2]), mn = make_int3(cmin[0], cmin[1], cmin[2]);
  cluster_init_kernel<<<g, 256, 0, s>>>(vox_bits, n_cov, parent);
  cluster_box_init_kernel<<<g, 256, 0, s>>>(n_cov, size, box);
  cluster_union_kernel<<<g, 256, 0, s>>>(vox_bits, n_cov, (const long long*)keys, dv, offs, noffs, parent);
  cluster_label_kernel<<<g, 256, 0, s>>>(n_cov, (const long long*)keys, mn, dv, parent, root, size, box);
}

}  // namespace asscpp

// include/asscpp_gpu.h: no cudaMalloc / cudaFree / allocating routine on this path, nothing synchronises; the three
// launches go on the caller's stream and leave ids + count in the caller's device buffers.
extern "C" int asscpp_uncovered_ids_noalloc(const uint32_t* d_voxel_bits, int64_t n_voxels, int32_t* d_ids, int64_t cap,
                                            int64_t* d_count, void* stream, char* err, int64_t err_len) {
  using namespace asscpp;
  if (err && err_len > 0) err[0] = 0;
  if (n_voxels < 0 || cap < 0 || !d_count || (n_voxels > 0 && !d_voxel_bits) || (cap > 0 && !d_ids)) {
    noalloc_error(err, err_len, "asscpp_uncovered_ids_noalloc: bad argument (n_voxels %lld, cap %lld)%s", n_voxels, cap, "");
    return ASSCPP_EINVAL;
  }
  if (n_voxels > (int64_t)ASSCPP_NOALLOC_MAX_VOXELS) {
    noalloc_error(err, err_len, "asscpp_uncovered_ids_noalloc: n_voxels %lld exceeds ASSCPP_NOALLOC_MAX_VOXELS %lld%s", n_voxels,
                  (long long)ASSCPP_NOALLOC_MAX_VOXELS, "");
    return ASSCPP_ERANGE;
  }
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kNoallocMaxDevices) {
    noalloc_error(err, err_len, "asscpp_uncovered_ids_noalloc: cudaGetDevice -> %lld (device %lld) %s", (long long)e, dev,
                  e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range");
    return e != cudaSuccess ? ASSCPP_ECUDA : ASSCPP_ERANGE;
  }
  // set-up, once per device, with that device current: the device address of the scratch array
  int32_t* scratch = g_noalloc_ptr[dev].load(std::memory_order_acquire);
  if (!scratch) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      noalloc_error(err, err_len, "asscpp_uncovered_ids_noalloc: cudaGetSymbolAddress -> %lld (device %lld) %s", (long long)e, dev,
                    cudaGetErrorString(e));
      return ASSCPP_ECUDA;
    }
    scratch = (int32_t*)p;
    g_noalloc_ptr[dev].store(scratch, std::memory_order_release);
  }
  // every word of the scratch the emit kernel reads is rewritten by the count + scan kernels of this call and
  // *d_count is written by the scan kernel (or the memset below): nothing has to start at zero
  const int32_t zero3[3] = {0, 0, 0}, one3[3] = {1, 1, 1};
  launch_uncovered(d_voxel_bits, n_voxels, scratch, d_count, nullptr, zero3, one3, 0.0f, d_ids, nullptr, cap,
                   (cudaStream_t)stream);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    noalloc_error(err, err_len, "asscpp_uncovered_ids_noalloc: launch -> %lld (device %lld) %s", (long long)e, dev,
                  cudaGetErrorString(e));
    return ASSCPP_ECUDA;
  }
  return ASSCPP_OK;
}
