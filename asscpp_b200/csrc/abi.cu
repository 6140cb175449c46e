// abi.cu -- the C ABI of include/asscpp_gpu.h: context, device memory, kernel sequencing.
// No torch, no CPU fallback: every compute entry point launches the sm_100a kernels of kernels.cu.
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "common.h"
#include "kernels.h"
#include "preprocess.h"

using namespace asscpp;

namespace {

std::string g_create_error;

// ---- NCCL through dlopen (the library is torch's bundled libnccl.so.2 or the system one) -------
struct NcclUniqueId { char internal[128]; };
typedef void* NcclComm;
struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(NcclUniqueId*) = nullptr;
  int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
  int (*CommInitAll)(NcclComm*, int, const int*) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, NcclComm, cudaStream_t) = nullptr;
  int (*CommDestroy)(NcclComm) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool load(std::string* err) {
    if (handle) return true;
    const char* env = getenv("ASSCPP_NCCL_LIB");
    const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      if (!nm || !*nm) continue;
      handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (handle) break;
    }
    if (!handle) { *err = std::string("cannot dlopen libnccl: ") + dlerror(); return false; }
    GetUniqueId = (int (*)(NcclUniqueId*))dlsym(handle, "ncclGetUniqueId");
    CommInitRank = (int (*)(NcclComm*, int, NcclUniqueId, int))dlsym(handle, "ncclCommInitRank");
    AllGather = (int (*)(const void*, void*, size_t, int, NcclComm, cudaStream_t))dlsym(handle, "ncclAllGather");
    CommDestroy = (int (*)(NcclComm))dlsym(handle, "ncclCommDestroy");
    CommInitAll = (int (*)(NcclComm*, int, const int*))dlsym(handle, "ncclCommInitAll");
    GetErrorString = (const char* (*)(int))dlsym(handle, "ncclGetErrorString");
    if (!GetUniqueId || !CommInitRank || !AllGather || !CommDestroy) { *err = "libnccl lacks required symbols"; return false; }
    return true;
  }
};
NcclApi g_nccl;
constexpr int kNcclUint8 = 1;

struct CoveredSet {
  uint32_t* buf = nullptr;  // [pts words | vox words] contiguous (one allgather moves both)
  bool used = false;
};

}  // namespace

struct AsscppCtx {
  int device = 0;
  int sm_count = 0;
  cudaStream_t stream = nullptr;
  AsscppParams params;
  FrustumConst fc;
  ModelDev m;
  std::vector<void*> allocs;  // everything cudaMalloc'ed for the model + scratch
  int64_t cov_keys_n = 0;
  int64_t* d_cov_keys = nullptr;
  int32_t cov_min_b[3], cov_div[3];
  // sizes
  int64_t pts_words = 0, vox_words = 0;  // uint32 words, even counts
  int ctas = 0, ordered_ctas = 8;
  int cta_threads = 1024;   // eval_views_kernel: 1024 -> one CTA per SM, 512 -> two (half the shared-memory grid each)
  size_t eval_smem_bytes_ = 0;
  size_t smem_bytes = 0;
  uint32_t smem_grid_words = 0, eval_grid_words = 0;
  int64_t chunk_cap = 0, list_cap = 0, grid_cap_words = 0, cov_words = 0;
  uint2* d_chunk_list = nullptr;
  uint32_t *d_vox_list = nullptr, *d_vis_list = nullptr, *d_grid_scratch = nullptr, *d_cov_scratch = nullptr;
  int64_t vis_cap = 0;
  uint32_t* d_view_cov = nullptr;      // split-group mode: per-view coverage bitsets
  int64_t view_cov_cap = 0;
  int64_t split_group_max_views = 4096;
  // small batches (at most kSmallMaxViews views, host poses, no accumulator, no communicator): ONE launch per evaluation
  // (eval_views_small_kernel); gains and counters are mirrored into mapped host memory, so reading the gains back is a
  // stream synchronisation and nothing else.  A view that needs the ordered replay (quirk Q5) sends the whole batch
  // through the general path when its results are asked for (finalize_small).
  struct SmallHost {
    uint2 gains[kSmallMaxViews];
    int32_t counters[4];
    ViewTrig trig[kSmallMaxViews];
    AsscppViewResult results[kSmallMaxViews];
    AccRaw acc[kSmallMaxViews];
  };
  SmallHost* h_small = nullptr;  // cudaHostAlloc, mapped
  SmallHost* d_small = nullptr;  // the same memory as the device sees it
  uint32_t* d_group_ticket = nullptr;
  bool small_enabled = true, small_pending = false, small_valid = false, counters_clean = false;
  std::vector<double> small_poses;
  int small_covered = -1, small_group = 1;
  uint32_t small_flags = 0;
  int64_t small_launches = 0, small_fallbacks = 0;
  int32_t small_epoch = 0;
  int small_spin_us = 200;   // finalize_small spins this long on the completion flag before it synchronises the stream
  uint4* d_ray_buf = nullptr;
  int64_t ray_cap = 0;
  size_t l2_persist_bytes = 0;  // persisting-L2 window over the ray queues (0: not set)
  size_t l2_persist_max = 0;
  int chunk_cull = 0;  // coarse lattice (>= 16 points per occupied voxel): cull boundary chunks chunk by chunk
  int walk_thresh = 16;
  int setup_iters = 1, halo_iters = 3;
  uint8_t* d_vstate = nullptr;
  int32_t* d_counters = nullptr;  // [0] work counter, [1] ordered count
  unsigned long long* d_pop = nullptr;  // 2 popcount cells
  unsigned long long* d_phase = nullptr;  // per-phase cycle counters, allocated by asscpp_debug_phase_cycles
  // per-batch buffers (grown on demand)
  int64_t view_cap = 0;
  double* d_poses = nullptr;
  ViewDev* d_views = nullptr;
  AsscppViewResult* d_results = nullptr;
  uint2* d_gains = nullptr;     // packed {gain_pts, gain_vox} per view (what asscpp_get_gain / asscpp_allgather_gain move)
  std::vector<uint2> h_gains;
  int32_t* d_ordered_list = nullptr;
  ViewTrig* d_trig = nullptr;   // ASSCPP_EVAL_ACCURACY
  AccRaw* d_acc = nullptr;
  std::vector<ViewTrig> h_trig;
  std::vector<AccRaw> h_acc;
  double acc_scale_d = 1, acc_scale_z = 1;
  bool last_had_acc = false, last_had_cells = false;
  int64_t last_nviews = -1;
  std::vector<AsscppViewResult> h_results;
  // uncovered regions (allocated on first use)
  int32_t* d_region = nullptr;
  int32_t* d_region_ids = nullptr;
  float* d_region_ctr = nullptr;
  int64_t* d_region_total = nullptr;
  int3* d_region_offs = nullptr;
  // staging
  uint32_t* d_tmp_bits = nullptr;  // pts_words, original-index space staging
  int64_t idx_cap = 0;
  int32_t* d_idx = nullptr;
  int64_t cloud_cap = 0;
  float* d_cloud = nullptr;
  std::vector<CoveredSet> sets;
  int tmp_set = -1;
  // timing
  cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
  bool timing = false;
  struct TimedSpan { int kind; cudaEvent_t e0, e1; };  // kind 0 eval_views_kernel, 1 allgather, 2 OR-allreduce
  std::vector<TimedSpan> kernel_events;
  size_t kernel_events_used = 0;
  float last_kernel_ms = 0;
  int64_t last_kernel_launches = 0;
  float last_coll_ms[2] = {0, 0};
  int64_t last_coll_calls[2] = {0, 0};
  int64_t launches = 0;
  // nccl
  NcclComm comm = nullptr;
  int nranks = 1, rank = 0;
  void* d_gather = nullptr;
  size_t gather_bytes = 0;
  int64_t* d_cnt = nullptr;  // [1 + nranks] equal-count check of the host-output allgathers
  std::string err;
};

namespace {

#define CU(ctx, expr)                                                                       \
  do {                                                                                      \
    cudaError_t e__ = (expr);                                                               \
    if (e__ != cudaSuccess) {                                                               \
      (ctx)->err = std::string(#expr) + ": " + cudaGetErrorString(e__);                     \
      return ASSCPP_ECUDA;                                                                  \
    }                                                                                       \
  } while (0)

template <class T>
int dev_alloc(AsscppCtx* c, T** p, int64_t count, bool track = true) {
  *p = nullptr;
  size_t bytes = (size_t)std::max<int64_t>(count, 1) * sizeof(T);
  cudaError_t e = cudaMalloc((void**)p, bytes);
  if (e != cudaSuccess) {
    c->err = std::string("cudaMalloc(") + std::to_string(bytes) + "): " + cudaGetErrorString(e);
    return e == cudaErrorMemoryAllocation ? ASSCPP_ENOMEM : ASSCPP_ECUDA;
  }
  if (track) c->allocs.push_back(*p);
  return ASSCPP_OK;
}

template <class T>
int upload(AsscppCtx* c, const T** dst, const std::vector<T>& src) {
  T* p;
  int rc = dev_alloc(c, &p, (int64_t)src.size());
  if (rc) return rc;
  if (!src.empty()) CU(c, cudaMemcpy(p, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice));
  *dst = p;
  return ASSCPP_OK;
}

int64_t set_words(const AsscppCtx* c) { return c->pts_words + c->vox_words; }

CoveredDev set_dev(const AsscppCtx* c, int id) {
  CoveredDev d;
  d.pts = c->sets[id].buf;
  d.vox = c->sets[id].buf + c->pts_words;
  return d;
}

bool valid_set(const AsscppCtx* c, int id) {
  return id >= 0 && id < (int)c->sets.size() && c->sets[id].used;
}

int new_set(AsscppCtx* c, int* out) {
  int id = -1;
  for (size_t i = 0; i < c->sets.size(); ++i)
    if (!c->sets[i].used) { id = (int)i; break; }
  if (id < 0) { c->sets.push_back(CoveredSet()); id = (int)c->sets.size() - 1; }
  if (!c->sets[id].buf) {
    int rc = dev_alloc(c, &c->sets[id].buf, set_words(c), false);
    if (rc) return rc;
  }
  CU(c, cudaMemsetAsync(c->sets[id].buf, 0, (size_t)set_words(c) * 4, c->stream));
  c->sets[id].used = true;
  *out = id;
  return ASSCPP_OK;
}

int ensure_views(AsscppCtx* c, int64_t nviews) {
  if (nviews <= c->view_cap) return ASSCPP_OK;
  CU(c, cudaStreamSynchronize(c->stream));
  int64_t cap = std::max<int64_t>(nviews, c->view_cap * 2);
  cudaFree(c->d_poses); cudaFree(c->d_views); cudaFree(c->d_results); cudaFree(c->d_ordered_list);
  cudaFree(c->d_trig); cudaFree(c->d_acc); cudaFree(c->d_gains);
  c->d_poses = nullptr; c->d_views = nullptr; c->d_results = nullptr; c->d_ordered_list = nullptr;
  c->d_trig = nullptr; c->d_acc = nullptr; c->d_gains = nullptr;
  c->view_cap = 0;
  int rc;
  if ((rc = dev_alloc(c, &c->d_gains, cap, false))) return rc;
  if ((rc = dev_alloc(c, &c->d_poses, cap * 7, false))) return rc;
  if ((rc = dev_alloc(c, &c->d_views, cap, false))) return rc;
  if ((rc = dev_alloc(c, &c->d_results, cap, false))) return rc;
  if ((rc = dev_alloc(c, &c->d_ordered_list, cap, false))) return rc;
  if ((rc = dev_alloc(c, &c->d_trig, cap, false))) return rc;
  if ((rc = dev_alloc(c, &c->d_acc, cap, false))) return rc;
  c->view_cap = cap;
  return ASSCPP_OK;
}

// between asscpp_timer_begin / _end: bracket a kernel or a collective with events on the context's stream
AsscppCtx::TimedSpan* span_begin(AsscppCtx* c, int kind) {
  if (!c->timing) return nullptr;
  if (c->kernel_events_used == c->kernel_events.size()) {
    AsscppCtx::TimedSpan t;
    t.kind = kind;
    if (cudaEventCreate(&t.e0) != cudaSuccess || cudaEventCreate(&t.e1) != cudaSuccess) return nullptr;
    c->kernel_events.push_back(t);
  }
  AsscppCtx::TimedSpan* t = &c->kernel_events[c->kernel_events_used++];
  t->kind = kind;
  cudaEventRecord(t->e0, c->stream);
  return t;
}
void span_end(AsscppCtx* c, AsscppCtx::TimedSpan* t) {
  if (t) cudaEventRecord(t->e1, c->stream);
}

bool takes_small_path(const AsscppCtx* c, int64_t nviews, int accum_set, int group, const double* h_poses) {
  return c->small_enabled && c->h_small && h_poses && !c->comm && nviews > 0 && nviews <= kSmallMaxViews && accum_set < 0 &&
         (group == 1 || (group > 1 && group <= kSmallMaxGroup && nviews <= c->split_group_max_views));
}

// planes -> eval -> ordered, all on the context's stream
int run_eval(AsscppCtx* c, const double* d_poses, int64_t nviews, int covered_set, int accum_set,
             int64_t accum_min_visible, uint32_t flags, const double* h_poses = nullptr, int group = 1,
             bool allow_small = true) {
  if (nviews < 0 || nviews > 0x7fffffffLL) { c->err = "nviews out of range"; return ASSCPP_EINVAL; }
  if (group < 1) { c->err = "group size must be at least 1"; return ASSCPP_EINVAL; }
  if (covered_set != -1 && !valid_set(c, covered_set)) { c->err = "bad covered set handle"; return ASSCPP_EINVAL; }
  if (accum_set != -1 && !valid_set(c, accum_set)) { c->err = "bad accumulator set handle"; return ASSCPP_EINVAL; }
  if (accum_set != -1 && accum_set == covered_set) {
    c->err = "the accumulator must differ from the set scored against (gains would depend on the view order)";
    return ASSCPP_EINVAL;
  }
  c->last_nviews = nviews;
  c->last_had_acc = (flags & ASSCPP_EVAL_ACCURACY) != 0;
  const bool want_cells = (flags & ASSCPP_EVAL_VIEW_CELLS) != 0;
  if (want_cells && (accum_set >= 0 || nviews > c->split_group_max_views || group > kSmallMaxGroup)) {
    c->err = "ASSCPP_EVAL_VIEW_CELLS: at most 4096 views per call, groups of at most 8, no accumulator set";
    return ASSCPP_EINVAL;
  }
  c->last_had_cells = want_cells;
  c->small_pending = false; c->small_valid = false;
  if (nviews == 0) return ASSCPP_OK;
  const bool small = allow_small && takes_small_path(c, nviews, accum_set, group, h_poses);
  if (small) {
    if (flags & ASSCPP_EVAL_ACCURACY) {
      CU(c, cudaStreamSynchronize(c->stream));  // the previous small batch may still be reading the mapped constants
      for (int64_t v = 0; v < nviews; ++v) viewport_trig_host(h_poses + 7 * v, &c->h_small->trig[(size_t)v]);
    }
  } else if (flags & ASSCPP_EVAL_ACCURACY) {
    if (h_poses) {
      // yaw / pitch and their sines and cosines with the host's libm, as the reference computes them
      // (occlusion_culling.cpp:556-584)
      CU(c, cudaStreamSynchronize(c->stream));  // h_trig may still be in flight from the previous batch
      c->h_trig.resize((size_t)nviews);
      for (int64_t v = 0; v < nviews; ++v) viewport_trig_host(h_poses + 7 * v, &c->h_trig[(size_t)v]);
      CU(c, cudaMemcpyAsync(c->d_trig, c->h_trig.data(), (size_t)nviews * sizeof(ViewTrig), cudaMemcpyHostToDevice,
                            c->stream));
    } else {
      launch_trig(d_poses, nviews, c->d_trig, c->stream);
      c->launches += 1;
    }
  }
  if (!small || !c->counters_clean) CU(c, cudaMemsetAsync(c->d_counters, 0, 4 * sizeof(int32_t), c->stream));
  c->counters_clean = small;  // the small kernel leaves them zeroed
  if (!small) launch_planes(d_poses, nviews, c->fc, c->d_views, c->stream);
  EvalArgs a;
  a.m = c->m;
  a.views = c->d_views;
  a.nviews = nviews;
  a.work_counter = c->d_counters;
  a.results = c->d_results;
  a.gains = c->d_gains;
  a.error_count = c->d_counters + 2;
  a.covered.pts = nullptr; a.covered.vox = nullptr;
  if (covered_set >= 0) a.covered = set_dev(c, covered_set);
  a.accum.pts = nullptr; a.accum.vox = nullptr;
  if (accum_set >= 0) a.accum = set_dev(c, accum_set);
  a.accum_min_visible = accum_min_visible;
  a.flags = flags;
  a.trig = small ? c->d_small->trig : c->d_trig;
  a.acc = (flags & ASSCPP_EVAL_ACCURACY) ? c->d_acc : nullptr;
  a.small = small ? 1 : 0;
  a.chunk_cull = c->chunk_cull;
  a.fc = c->fc;
  a.host_gains = small ? c->d_small->gains : nullptr;
  a.host_counters = small ? c->d_small->counters : nullptr;
  a.host_results = small ? c->d_small->results : nullptr;
  a.host_acc = small ? c->d_small->acc : nullptr;
  a.group_ticket = c->d_group_ticket;
  a.acc_scale_d = c->acc_scale_d; a.acc_scale_z = c->acc_scale_z;
  a.chunk_list = c->d_chunk_list; a.chunk_cap = c->chunk_cap;
  a.vox_list = c->d_vox_list; a.list_cap = c->list_cap;
  a.vis_list = c->d_vis_list; a.vis_cap = c->vis_cap;
  a.ray_buf = c->d_ray_buf; a.ray_cap = c->ray_cap;
  a.walk_thresh = c->walk_thresh;
  a.setup_iters = c->setup_iters; a.halo_iters = c->halo_iters;
  a.grid_scratch = c->d_grid_scratch; a.grid_cap_words = c->grid_cap_words;
  a.cov_scratch = c->d_cov_scratch; a.cov_words = c->cov_words;
  a.group = group; a.cov_direct = (c->m.cov_same && group == 1) ? 1 : 0;
  a.view_cov = nullptr; a.split_group = 0;
  // Small grouped batches (the A* use: the sensors of a few dozen children): a group per CTA leaves most SMs idle and
  // runs the group's views one after the other.  Split-group mode makes every VIEW a work unit; each writes the
  // coverage-voxel bitset of its visible cloud and group_gain_kernel turns those into the union gains.
  const int64_t split_max = c->split_group_max_views;
  if ((group > 1 || want_cells) && nviews <= split_max && accum_set < 0) {
    const int64_t need = nviews * c->cov_words;
    if (need > c->view_cov_cap) {
      CU(c, cudaStreamSynchronize(c->stream));
      cudaFree(c->d_view_cov);
      c->d_view_cov = nullptr; c->view_cov_cap = 0;
      int rc = dev_alloc(c, &c->d_view_cov, need, false);
      if (rc) return rc;
      c->view_cov_cap = need;
    }
    if (!small) CU(c, cudaMemsetAsync(c->d_view_cov, 0, (size_t)need * 4, c->stream));  // small: every CTA clears its view's row
    a.view_cov = c->d_view_cov; a.split_group = group; a.group = 1; a.cov_direct = 0;
  }
  a.vstate = c->d_vstate;
  a.ordered_list = c->d_ordered_list;
  a.ordered_count = c->d_counters + 1;
  a.smem_grid_words = c->smem_grid_words;
  a.phase_cycles = c->d_phase;
  int ctas = (int)std::min<int64_t>(c->ctas, (nviews + a.group - 1) / a.group);
  AsscppCtx::TimedSpan* evp = span_begin(c, 0);
  a.smem_grid_words = c->eval_grid_words;
  if (small) {
    ViewDev tiny[8];
    const bool tiny_batch = nviews <= std::min(8, tiny_max_views());
    for (int64_t v = 0; tiny_batch && v < nviews; ++v) {
      frustum_planes_host_fc(c->fc, h_poses + 7 * v, tiny[v].planes, tiny[v].origin, nullptr);
      tiny[v].pad = 0.0f;
    }
    a.epoch = ++c->small_epoch;
    launch_eval_small(a, h_poses, tiny_batch ? tiny : nullptr, ctas, c->cta_threads, c->eval_smem_bytes_, c->stream);
    span_end(c, evp);
    c->launches += 1;
    c->small_launches += 1;
    CU(c, cudaGetLastError());
    c->small_poses.assign(h_poses, h_poses + 7 * nviews);
    c->small_covered = covered_set; c->small_group = group; c->small_flags = flags;
    c->small_pending = true;
    return ASSCPP_OK;
  }
  launch_eval(a, ctas, c->cta_threads, c->eval_smem_bytes_, c->stream);
  a.smem_grid_words = c->smem_grid_words;
  span_end(c, evp);
  launch_ordered(a, c->ordered_ctas, c->smem_bytes, c->stream);
  c->launches += 3;
  if (a.view_cov) { launch_group_gain(a, a.split_group, c->stream); c->launches += 1; }
  CU(c, cudaGetLastError());
  return ASSCPP_OK;
}

// Before anything reads the results of the last evaluation: a small batch is complete once the stream has drained and
// no view asked for the ordered replay; otherwise the batch is evaluated again by the general path.
int finalize_small(AsscppCtx* c) {
  if (!c->small_pending) return ASSCPP_OK;
  {
    // the kernel's last CTA writes the launch's epoch into mapped host memory after everything else (system-scope
    // fences): seeing it is as good as the stream synchronisation and several microseconds sooner
    volatile int32_t* flag = &c->h_small->counters[2];
    bool seen = *flag == c->small_epoch;
    if (!seen && c->small_spin_us > 0) {
      const auto t0 = std::chrono::steady_clock::now();
      for (;;) {
        for (int k = 0; k < 64 && !seen; ++k) seen = *flag == c->small_epoch;
        if (seen || std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(c->small_spin_us)) break;
      }
    }
    if (!seen) CU(c, cudaStreamSynchronize(c->stream));
    std::atomic_thread_fence(std::memory_order_acquire);
  }
  c->small_pending = false;
  if (c->h_small->counters[0] == 0) { c->small_valid = true; return ASSCPP_OK; }
  c->small_fallbacks += 1;
  const int64_t n = (int64_t)(c->small_poses.size() / 7);
  std::vector<double> poses;
  poses.swap(c->small_poses);
  CU(c, cudaMemcpyAsync(c->d_poses, poses.data(), poses.size() * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  int rc = run_eval(c, c->d_poses, n, c->small_covered, -1, 0, c->small_flags, poses.data(), c->small_group, false);
  if (rc) return rc;
  CU(c, cudaStreamSynchronize(c->stream));  // poses (pageable) and the trig constants go out of scope
  return ASSCPP_OK;
}

int fetch_results(AsscppCtx* c, int64_t nviews) {
  if (c->last_nviews < 0) { c->err = "no evaluation has been issued"; return ASSCPP_ESTATE; }
  if (nviews > c->last_nviews) { c->err = "more results requested than views evaluated"; return ASSCPP_EINVAL; }
  if (int rc = finalize_small(c)) return rc;
  c->h_results.resize((size_t)std::max<int64_t>(nviews, 1));
  if (c->small_valid) {  // mirrored into mapped host memory by the kernel, which has finished
    if (nviews > 0) memcpy(c->h_results.data(), c->h_small->results, (size_t)nviews * sizeof(AsscppViewResult));
  } else {
    if (nviews > 0)
      CU(c, cudaMemcpyAsync(c->h_results.data(), c->d_results, (size_t)nviews * sizeof(AsscppViewResult),
                            cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
  }
  for (int64_t v = 0; v < nviews; ++v)
    if (c->h_results[v].status & kVsError) { c->err = "internal scratch bound violated for a view"; return ASSCPP_ERANGE; }
  return ASSCPP_OK;
}

}  // namespace

extern "C" {

int asscpp_abi_version(void) { return ASSCPP_ABI_VERSION; }

void asscpp_default_params(AsscppParams* p) {
  if (!p) return;
  p->leaf = 0.5f; p->hfov_deg = 58.0f; p->vfov_deg = 45.0f; p->near_dist = 0.7f; p->far_dist = 6.0f;
  p->pcd_origin[0] = p->pcd_origin[1] = p->pcd_origin[2] = 0.0f;
  p->cov_leaf = 0.5f;
}

const char* asscpp_last_error(const AsscppCtx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int asscpp_create(int device, const float* xyz, int64_t n, const AsscppParams* params, AsscppCtx** out) {
  if (!out) return ASSCPP_EINVAL;
  *out = nullptr;
  if ((!xyz && n > 0) || !params || n < 0 || n > 0x7fffffffLL || !(params->leaf > 0) || !(params->cov_leaf > 0) ||
      !(params->far_dist > params->near_dist) || !(params->near_dist > 0)) {
    g_create_error = "invalid arguments to asscpp_create";
    return ASSCPP_EINVAL;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) {
    cudaGetLastError();
    g_create_error = "no usable CUDA device (this library has no CPU path)";
    return ASSCPP_ENODEVICE;
  }
  AsscppCtx* c = new AsscppCtx();
  auto fail = [&](int rc) {
    g_create_error = c->err;
    asscpp_destroy(c);
    return rc;
  };
  c->device = device;
  c->params = *params;
  c->fc = frustum_const(*params);
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) { c->err = cudaGetErrorString(e); return fail(ASSCPP_ENODEVICE); }
  int major = 0, optin = 0;
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
  cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
  cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
  if (major < 10) { c->err = "device is not sm_100 class; kernels are built for sm_100a only"; return fail(ASSCPP_ENODEVICE); }
  if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) {
    c->err = cudaGetErrorString(e);
    return fail(ASSCPP_ECUDA);
  }
  cudaEventCreate(&c->ev_begin);
  cudaEventCreate(&c->ev_end);

  ModelHost h;
  std::string perr;
  int rc = preprocess_model(xyz, n, *params, &h, &perr);
  if (rc) { c->err = perr; return fail(rc); }
  ModelDev& m = c->m;
  memset(&m, 0, sizeof(m));
  m.n = h.n; m.n_vox = h.n_vox; m.n_cov = h.n_cov; m.n_chunks = h.n_chunks; m.n_super = h.n_super;
  m.leaf = params->leaf; m.inv_leaf = h.inv_leaf; m.half_leaf = params->leaf * 0.5f;
  for (int k = 0; k < 3; ++k) {
    m.pcd_origin[k] = params->pcd_origin[k];
    m.min_b[k] = h.min_b[k]; m.max_b[k] = h.max_b[k];
    c->cov_min_b[k] = h.cov_min_b[k]; c->cov_div[k] = h.cov_div[k];
  }
  m.dist_thr_sq = h.dist_thr_sq;
  m.max_abs_coord = 0;
  for (int k = 0; k < 3; ++k)
    m.max_abs_coord = std::max(m.max_abs_coord, std::max(std::fabs((float)h.min_b[k]), std::fabs((float)h.max_b[k] + 1.0f)) * params->leaf);
  m.max_abs_coord *= 1.001f;
  {
    // fixed-point scales of the accuracy sums: term * 2^S < 2^70 and (hi part) * visible < 2^63
    double dmax = std::sqrt((double)params->far_dist * params->far_dist + 0.25 * (double)c->fc.fp_w * c->fc.fp_w +
                            0.25 * (double)c->fc.fp_h * c->fc.fp_h) * 1.01 + 1e-3;
    double zmax = (double)m.max_abs_coord + 1e-3;
    int extra = 0;
    while (((int64_t)1 << (25 + extra)) < n) ++extra;
    auto scale_for = [&](double vmax) {
      double term = 0.0000285 * vmax * vmax;
      int e = 0;
      std::frexp(term, &e);  // term < 2^e
      return std::ldexp(1.0, 70 - e - extra);
    };
    c->acc_scale_d = scale_for(dmax);
    c->acc_scale_z = scale_for(zmax);
  }
  m.cov_same = h.cov_same;
  static_assert(sizeof(Float4) == sizeof(float4) && sizeof(Int4) == sizeof(int4), "layout");
  if ((rc = upload(c, (const Float4**)&m.pts, h.pts))) return fail(rc);
  if ((rc = upload(c, &m.orig, h.orig))) return fail(rc);
  if ((rc = upload(c, &m.pos_of, h.pos_of))) return fail(rc);
  if ((rc = upload(c, (const Int4**)&m.vox, h.vox))) return fail(rc);
  if ((rc = upload(c, &m.vox_cov, h.vox_cov))) return fail(rc);
  if ((rc = upload(c, &m.pt_cov, h.pt_cov))) return fail(rc);
  if ((rc = upload(c, (const Float4**)&m.chunk_lo, h.chunk_lo))) return fail(rc);
  if ((rc = upload(c, (const Float4**)&m.chunk_hi, h.chunk_hi))) return fail(rc);
  if ((rc = upload(c, &m.chunk_vox, h.chunk_vox))) return fail(rc);
  if ((rc = upload(c, (const Float4**)&m.super_lo, h.super_lo))) return fail(rc);
  if ((rc = upload(c, (const Float4**)&m.super_hi, h.super_hi))) return fail(rc);
  {
    const int64_t* dk;
    if ((rc = upload(c, &dk, h.cov_keys))) return fail(rc);
    c->d_cov_keys = const_cast<int64_t*>(dk);
    c->cov_keys_n = (int64_t)h.cov_keys.size();
  }
  c->pts_words = 2 * ((m.n + 63) / 64);
  c->vox_words = 2 * ((m.n_cov + 63) / 64);
  if (c->pts_words == 0) c->pts_words = 2;
  if (c->vox_words == 0) c->vox_words = 2;

  // ---- launch geometry + scratch ----------------------------------------------------------------
  if (const char* ct = getenv("ASSCPP_CTA_THREADS")) {
    int v = atoi(ct);
    if (v == 256 || v == 512 || v == 1024) c->cta_threads = v;
  }
  const int per_sm = 1024 / c->cta_threads;
  c->ctas = c->sm_count * per_sm;
  size_t hdr = eval_smem_header_bytes();
  size_t avail = (size_t)optin > hdr + 64 ? (size_t)optin - hdr - 64 : 0;
  c->smem_grid_words = (uint32_t)(avail / 4);
  c->smem_bytes = eval_smem_bytes(c->smem_grid_words);  // ordered_views_kernel (always 1024 threads)
  {
    // per-CTA share of the SM's shared memory (1 KB per CTA is reserved by the system)
    size_t sm_total = 0;
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerMultiprocessor, device);
    sm_total = (size_t)v;
    size_t share = per_sm == 1 ? (size_t)optin : std::min<size_t>((size_t)optin, sm_total / per_sm - 1024);
    size_t av = share > hdr + 64 ? share - hdr - 64 : 0;
    c->eval_grid_words = (uint32_t)(av / 4);
    c->eval_smem_bytes_ = eval_smem_bytes(c->eval_grid_words);
  }
  if ((e = configure_kernels(c->smem_bytes)) != cudaSuccess) { c->err = cudaGetErrorString(e); return fail(ASSCPP_ECUDA); }
  double ext = frustum_max_extent(*params);
  double cells_cap = 1;
  for (int k = 0; k < 3; ++k) {
    double model_div = (double)h.max_b[k] - h.min_b[k] + 1;
    double a = std::min(model_div, std::floor(ext * (double)h.inv_leaf) + 4.0) + 4.0;  // + halo and sentinel layers
    cells_cap *= std::max(a, 1.0);
  }
  if (cells_cap > 2147483647.0) cells_cap = 2147483647.0;
  c->grid_cap_words = (int64_t)(cells_cap / 32.0) + 2;
  // Coarse lattices: no view's grid can be larger than the frustum's cell capacity, so the kernels ask for no more
  // shared memory than that (the rest of the SM's 256 KB stays L1 for the voxel table and the point runs).
  // ASSCPP_FULL_SMEM=1 keeps the full request (A/B).
  if (!getenv("ASSCPP_FULL_SMEM")) {
    const int64_t need = ((c->grid_cap_words + 63) / 64) * 64;
    if (need < (int64_t)c->eval_grid_words) {
      c->eval_grid_words = (uint32_t)need;
      c->eval_smem_bytes_ = eval_smem_bytes(c->eval_grid_words);
    }
    if (need < (int64_t)c->smem_grid_words) {
      c->smem_grid_words = (uint32_t)need;
      c->smem_bytes = eval_smem_bytes(c->smem_grid_words);
    }
  }
  c->chunk_cap = std::max<int64_t>(m.n_chunks, 1);
  c->list_cap = m.n_vox + m.n / 32 + m.n_chunks + 64;
  c->cov_words = c->vox_words;
  if ((rc = dev_alloc(c, &c->d_chunk_list, c->chunk_cap * c->ctas))) return fail(rc);
  if ((rc = dev_alloc(c, &c->d_vox_list, c->list_cap * c->ctas))) return fail(rc);
  c->vis_cap = m.n_vox + 32;
  if ((rc = dev_alloc(c, &c->d_vis_list, c->vis_cap * c->ctas))) return fail(rc);
  c->ray_cap = ray_queue_cap();  // records per warp queue; 32 warps per CTA
  if ((rc = dev_alloc(c, &c->d_ray_buf, 3 * c->ray_cap * 32 * c->ctas))) return fail(rc);
  // The warps' ray queues are the most rewritten scratch (48 bytes written and read back per ray): ask L2 to keep
  // them (persisting access window over the one contiguous buffer) so that the stream of model points does not push
  // them out to DRAM.  ASSCPP_L2_PERSIST=0 turns it off; failure to set it up is not an error.
  {
    const char* pe = getenv("ASSCPP_L2_PERSIST");
    int max_persist = 0;
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, device);
    const size_t ray_bytes = (size_t)3 * c->ray_cap * 32 * c->ctas * sizeof(uint4);
    if ((!pe || atoi(pe) != 0) && max_persist > 0) {
      // (+ room for evict-last lines of the voxel lists when built with -DASSCPP_KEEP_LISTS: not adopted, kernels.cu)
      size_t extra = 0;
      if (const char* xe = getenv("ASSCPP_L2_PERSIST_EXTRA_MB")) extra = (size_t)std::max(0, atoi(xe)) << 20;
      const size_t want = std::min<size_t>((size_t)max_persist, ray_bytes + extra);
      int max_window = 0;
      cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, device);
      if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) {
        cudaStreamAttrValue attr;
        memset(&attr, 0, sizeof(attr));
        attr.accessPolicyWindow.base_ptr = c->d_ray_buf;
        attr.accessPolicyWindow.num_bytes = std::min<size_t>(ray_bytes, (size_t)std::max(max_window, 0));
        attr.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)want / (double)std::max<size_t>(ray_bytes, 1));
        c->l2_persist_max = (size_t)max_persist;
        attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
        if (cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &attr) == cudaSuccess) c->l2_persist_bytes = want;
      }
      cudaGetLastError();
    }
  }
  c->chunk_cull = (m.n_vox > 0 && m.n >= 16 * m.n_vox) ? 1 : 0;
  if (const char* e = getenv("ASSCPP_CHUNK_CULL")) c->chunk_cull = atoi(e) != 0;
  if (const char* e = getenv("ASSCPP_HALO_ITERS")) c->halo_iters = std::max(0, std::min(64, atoi(e)));
  if (const char* e = getenv("ASSCPP_SETUP_ITERS")) c->setup_iters = std::max(1, std::min(64, atoi(e)));
  if (const char* e = getenv("ASSCPP_SPLIT_GROUP_MAX")) c->split_group_max_views = std::max(0, atoi(e));
  if (const char* wt = getenv("ASSCPP_WALK_THRESH")) c->walk_thresh = std::max(1, std::min(32, atoi(wt)));
  if ((rc = dev_alloc(c, &c->d_grid_scratch, c->grid_cap_words * c->ctas))) return fail(rc);
  if ((rc = dev_alloc(c, &c->d_cov_scratch, c->cov_words * c->ctas))) return fail(rc);
  if ((rc = dev_alloc(c, &c->d_vstate, std::max<int64_t>(m.n_vox, 1) * c->ordered_ctas))) return fail(rc);
  if ((rc = dev_alloc(c, &c->d_counters, 4))) return fail(rc);
  cudaMemsetAsync(c->d_counters, 0, 4 * sizeof(int32_t), c->stream);
  c->counters_clean = true;
  if ((rc = dev_alloc(c, &c->d_group_ticket, kSmallMaxViews))) return fail(rc);
  cudaMemsetAsync(c->d_group_ticket, 0, kSmallMaxViews * sizeof(uint32_t), c->stream);
  if (getenv("ASSCPP_NO_SMALL")) c->small_enabled = false;
  if (const char* sp = getenv("ASSCPP_SMALL_SPIN_US")) c->small_spin_us = std::max(0, atoi(sp));
  {
    void* hp = nullptr;
    void* dp = nullptr;
    if (cudaHostAlloc(&hp, sizeof(AsscppCtx::SmallHost), cudaHostAllocMapped) == cudaSuccess &&
        cudaHostGetDevicePointer(&dp, hp, 0) == cudaSuccess) {
      memset(hp, 0, sizeof(AsscppCtx::SmallHost));
      c->h_small = (AsscppCtx::SmallHost*)hp;
      c->d_small = (AsscppCtx::SmallHost*)dp;
    } else {
      if (hp) cudaFreeHost(hp);
      cudaGetLastError();  // no mapped host memory: the general path serves small batches too
    }
  }
  if ((rc = dev_alloc(c, &c->d_pop, 2))) return fail(rc);
  if ((rc = dev_alloc(c, &c->d_tmp_bits, c->pts_words))) return fail(rc);
  cudaMemsetAsync(c->d_cov_scratch, 0, (size_t)c->cov_words * c->ctas * 4, c->stream);
  launch_fill_bytes(c->d_vstate, std::max<int64_t>(m.n_vox, 1) * c->ordered_ctas, 0xff, c->stream);
  c->launches += 1;
  if ((rc = new_set(c, &c->tmp_set))) return fail(rc);
  if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) { c->err = cudaGetErrorString(e); return fail(ASSCPP_ECUDA); }
  *out = c;
  return ASSCPP_OK;
}

int asscpp_destroy(AsscppCtx* c) {
  if (!c) return ASSCPP_OK;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
  if (c->h_small) cudaFreeHost(c->h_small);
  for (void* p : c->allocs) cudaFree(p);
  for (auto& s : c->sets) cudaFree(s.buf);
  cudaFree(c->d_poses); cudaFree(c->d_views); cudaFree(c->d_results); cudaFree(c->d_ordered_list);
  cudaFree(c->d_trig); cudaFree(c->d_acc); cudaFree(c->d_gains);
  cudaFree(c->d_idx); cudaFree(c->d_cloud); cudaFree(c->d_gather); cudaFree(c->d_cnt); cudaFree(c->d_view_cov);
  for (auto& ev : c->kernel_events) { cudaEventDestroy(ev.e0); cudaEventDestroy(ev.e1); }
  if (c->ev_begin) cudaEventDestroy(c->ev_begin);
  if (c->ev_end) cudaEventDestroy(c->ev_end);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
  return ASSCPP_OK;
}

int asscpp_model_info(const AsscppCtx* c, AsscppModelInfo* out) {
  if (!c || !out) return ASSCPP_EINVAL;
  out->n_points = c->m.n; out->n_occ_voxels = c->m.n_vox; out->n_cov_voxels = c->m.n_cov; out->n_chunks = c->m.n_chunks;
  for (int k = 0; k < 3; ++k) { out->min_b[k] = c->m.min_b[k]; out->max_b[k] = c->m.max_b[k]; }
  out->sm_count = c->sm_count;
  out->reserved = 0;
  return ASSCPP_OK;
}

int asscpp_eval_batch_device(AsscppCtx* c, const double* d_poses, int64_t nviews, int covered_set, uint32_t flags,
                             const AsscppViewResult** d_results) {
  if (!c || (!d_poses && nviews > 0)) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = ensure_views(c, nviews);
  if (rc) return rc;
  rc = run_eval(c, d_poses, nviews, covered_set, -1, 0, flags);
  if (rc) return rc;
  if (d_results) *d_results = c->d_results;
  return ASSCPP_OK;
}

int asscpp_eval_batch(AsscppCtx* c, const double* poses, int64_t nviews, int covered_set, uint32_t flags) {
  if (!c || (!poses && nviews > 0) || nviews < 0) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = ensure_views(c, nviews);
  if (rc) return rc;
  if (nviews > 0 && !takes_small_path(c, nviews, -1, 1, poses))
    CU(c, cudaMemcpyAsync(c->d_poses, poses, (size_t)nviews * 7 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  return run_eval(c, c->d_poses, nviews, covered_set, -1, 0, flags, poses);
}

int asscpp_eval_groups(AsscppCtx* c, const double* poses, int64_t nviews, int group_size, int covered_set,
                       uint32_t flags) {
  if (!c || (!poses && nviews > 0) || nviews < 0) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = ensure_views(c, nviews);
  if (rc) return rc;
  if (nviews > 0 && !takes_small_path(c, nviews, -1, group_size, poses))
    CU(c, cudaMemcpyAsync(c->d_poses, poses, (size_t)nviews * 7 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  return run_eval(c, c->d_poses, nviews, covered_set, -1, 0, flags, poses, group_size);
}

int asscpp_get_results(AsscppCtx* c, AsscppViewResult* out, int64_t nviews) {
  if (!c || (!out && nviews > 0) || nviews < 0) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = fetch_results(c, nviews);
  if (rc) return rc;
  if (nviews > 0) memcpy(out, c->h_results.data(), (size_t)nviews * sizeof(AsscppViewResult));
  return ASSCPP_OK;
}

int64_t asscpp_view_cells_words(const AsscppCtx* c) { return c ? c->cov_words : 0; }

int asscpp_get_view_cells(AsscppCtx* c, uint32_t* rows, int64_t nviews) {
  if (!c || (!rows && nviews > 0) || nviews < 0) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  if (c->last_nviews < 0 || !c->last_had_cells) { c->err = "the last evaluation was not issued with ASSCPP_EVAL_VIEW_CELLS"; return ASSCPP_ESTATE; }
  if (nviews > c->last_nviews) { c->err = "more rows requested than views evaluated"; return ASSCPP_EINVAL; }
  if (int rc = finalize_small(c)) return rc;
  if (nviews > 0)
    CU(c, cudaMemcpyAsync(rows, c->d_view_cov, (size_t)nviews * c->cov_words * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return ASSCPP_OK;
}

int asscpp_get_accuracy(AsscppCtx* c, AsscppViewAccuracy* out, int64_t nviews) {
  if (!c || (!out && nviews > 0) || nviews < 0) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  if (c->last_nviews < 0 || !c->last_had_acc) { c->err = "the last evaluation was not issued with ASSCPP_EVAL_ACCURACY"; return ASSCPP_ESTATE; }
  if (nviews > c->last_nviews) { c->err = "more results requested than views evaluated"; return ASSCPP_EINVAL; }
  if (int rc = finalize_small(c)) return rc;
  c->h_acc.resize((size_t)std::max<int64_t>(nviews, 1));
  if (c->small_valid) {
    if (nviews > 0) memcpy(c->h_acc.data(), c->h_small->acc, (size_t)nviews * sizeof(AccRaw));
  } else {
    if (nviews > 0)
      CU(c, cudaMemcpyAsync(c->h_acc.data(), c->d_acc, (size_t)nviews * sizeof(AccRaw), cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
  }
  auto decode = [](unsigned long long hi, unsigned long long lo, double scale) {
    unsigned __int128 t = ((unsigned __int128)hi << 32) + lo;  // exact sum of the scaled terms
    return (double)t / scale;                                  // one rounding; the division by 2^S is exact
  };
  auto unkey = [](int32_t k) {
    int32_t b = k >= 0 ? k : k ^ 0x7fffffff;
    float f;
    memcpy(&f, &b, 4);
    return f;
  };
  for (int64_t v = 0; v < nviews; ++v) {
    const AccRaw& r = c->h_acc[(size_t)v];
    out[v].err_sum_depth = decode(r.d_hi, r.d_lo, c->acc_scale_d);
    out[v].err_sum_z = decode(r.z_hi, r.z_lo, c->acc_scale_z);
    out[v].depth_min = unkey(r.min_key);
    out[v].depth_max = unkey(r.max_key);
    out[v].visible = r.visible;
    out[v].reserved = 0;
  }
  return ASSCPP_OK;
}

int asscpp_get_counts(AsscppCtx* c, uint32_t* survivors, uint32_t* visible, uint32_t* carry, int64_t nviews) {
  if (!c || nviews < 0) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = fetch_results(c, nviews);
  if (rc) return rc;
  for (int64_t v = 0; v < nviews; ++v) {
    if (survivors) survivors[v] = c->h_results[v].survivors;
    if (visible) visible[v] = c->h_results[v].visible;
    if (carry) carry[v] = c->h_results[v].carry;
  }
  return ASSCPP_OK;
}

// the packed 8 bytes per view only; the 64-byte records stay on the device
static int fetch_gains(AsscppCtx* c, const uint2* d_src, int64_t count) {
  c->h_gains.resize((size_t)std::max<int64_t>(count, 1));
  int32_t counters[4] = {0, 0, 0, 0};
  if (count > 0)
    CU(c, cudaMemcpyAsync(c->h_gains.data(), d_src, (size_t)count * sizeof(uint2), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemcpyAsync(counters, c->d_counters, sizeof(counters), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  if (counters[2]) { c->err = "internal scratch bound violated for a view"; return ASSCPP_ERANGE; }
  return ASSCPP_OK;
}

int asscpp_get_gain(AsscppCtx* c, uint32_t* gain_pts, uint32_t* gain_vox, int64_t nviews) {
  if (!c || nviews < 0) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  if (c->last_nviews < 0) { c->err = "no evaluation has been issued"; return ASSCPP_ESTATE; }
  if (nviews > c->last_nviews) { c->err = "more results requested than views evaluated"; return ASSCPP_EINVAL; }
  int rc = finalize_small(c);
  if (rc) return rc;
  if (c->small_valid) {
    // the kernel wrote the gains and the counters into mapped host memory: nothing to copy
    if (c->h_small->counters[1]) { c->err = "internal scratch bound violated for a view"; return ASSCPP_ERANGE; }
    for (int64_t v = 0; v < nviews; ++v) {
      if (gain_pts) gain_pts[v] = c->h_small->gains[(size_t)v].x;
      if (gain_vox) gain_vox[v] = c->h_small->gains[(size_t)v].y;
    }
    return ASSCPP_OK;
  }
  rc = fetch_gains(c, c->d_gains, nviews);
  if (rc) return rc;
  for (int64_t v = 0; v < nviews; ++v) {
    if (gain_pts) gain_pts[v] = c->h_gains[(size_t)v].x;
    if (gain_vox) gain_vox[v] = c->h_gains[(size_t)v].y;
  }
  return ASSCPP_OK;
}

int asscpp_eval_ex(AsscppCtx* c, const double* poses, int64_t nviews, const AsscppEvalOpts* o) {
  if (!c || !o || (!poses && nviews > 0) || nviews < 0) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = ensure_views(c, nviews);
  if (rc) return rc;
  const int group = o->group_size > 0 ? o->group_size : 1;
  if (o->poses_on_device)
    return run_eval(c, poses, nviews, o->covered_set, o->accum_set, o->accum_min_visible, o->flags, nullptr, group);
  if (nviews > 0 && !takes_small_path(c, nviews, o->accum_set, group, poses))
    CU(c, cudaMemcpyAsync(c->d_poses, poses, (size_t)nviews * 7 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  return run_eval(c, c->d_poses, nviews, o->covered_set, o->accum_set, o->accum_min_visible, o->flags, poses, group);
}

// ---- covered sets ---------------------------------------------------------------------------------
int asscpp_covered_new(AsscppCtx* c, int* set_out) {
  if (!c || !set_out) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  return new_set(c, set_out);
}

int asscpp_covered_free(AsscppCtx* c, int set) {
  if (!c || !valid_set(c, set) || set == c->tmp_set) return ASSCPP_EINVAL;
  c->sets[set].used = false;  // buffer is kept for reuse
  return ASSCPP_OK;
}

int asscpp_covered_clear(AsscppCtx* c, int set) {
  if (!c || !valid_set(c, set)) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  CU(c, cudaMemsetAsync(c->sets[set].buf, 0, (size_t)set_words(c) * 4, c->stream));
  return ASSCPP_OK;
}

int asscpp_covered_clone(AsscppCtx* c, int src, int* set_out) {
  if (!c || !set_out || !valid_set(c, src)) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = new_set(c, set_out);
  if (rc) return rc;
  CU(c, cudaMemcpyAsync(c->sets[*set_out].buf, c->sets[src].buf, (size_t)set_words(c) * 4, cudaMemcpyDeviceToDevice,
                        c->stream));
  return ASSCPP_OK;
}

int asscpp_covered_or_set(AsscppCtx* c, int dst, int src) {
  if (!c || !valid_set(c, dst) || !valid_set(c, src)) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  launch_or_words(c->sets[dst].buf, c->sets[src].buf, set_words(c), c->stream);
  c->launches += 1;
  CU(c, cudaGetLastError());
  return ASSCPP_OK;
}

int asscpp_covered_or_views(AsscppCtx* c, int set, const double* poses, int64_t nviews, int64_t min_visible) {
  if (!c || !valid_set(c, set) || (!poses && nviews > 0) || nviews < 0) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = ensure_views(c, nviews);
  if (rc) return rc;
  if (nviews > 0)
    CU(c, cudaMemcpyAsync(c->d_poses, poses, (size_t)nviews * 7 * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  return run_eval(c, c->d_poses, nviews, -1, set, min_visible, 0);
}

int asscpp_covered_or_points(AsscppCtx* c, int set, const int32_t* idx, int64_t n) {
  if (!c || !valid_set(c, set) || (!idx && n > 0) || n < 0) return ASSCPP_EINVAL;
  if (n == 0) return ASSCPP_OK;
  cudaSetDevice(c->device);
  if (n > c->idx_cap) {
    CU(c, cudaStreamSynchronize(c->stream));
    cudaFree(c->d_idx);
    c->d_idx = nullptr; c->idx_cap = 0;
    int rc = dev_alloc(c, &c->d_idx, n, false);
    if (rc) return rc;
    c->idx_cap = n;
  }
  CU(c, cudaMemcpyAsync(c->d_idx, idx, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
  launch_or_points(set_dev(c, set), c->d_idx, n, c->m, c->stream);
  c->launches += 1;
  CU(c, cudaGetLastError());
  CU(c, cudaStreamSynchronize(c->stream));  // idx may be pageable and reused by the caller
  return ASSCPP_OK;
}

int asscpp_covered_popcount(AsscppCtx* c, int set, int64_t* n_points, int64_t* n_voxels) {
  if (!c || !valid_set(c, set)) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  CoveredDev d = set_dev(c, set);
  CU(c, cudaMemsetAsync(c->d_pop, 0, 2 * sizeof(unsigned long long), c->stream));
  launch_popcount(d.pts, c->pts_words, c->d_pop, c->stream);
  launch_popcount(d.vox, c->vox_words, c->d_pop + 1, c->stream);
  c->launches += 2;
  unsigned long long h[2];
  CU(c, cudaMemcpyAsync(h, c->d_pop, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  if (n_points) *n_points = (int64_t)h[0];
  if (n_voxels) *n_voxels = (int64_t)h[1];
  return ASSCPP_OK;
}

int asscpp_covered_export(AsscppCtx* c, int set, uint64_t* point_bits, uint64_t* voxel_bits) {
  if (!c || !valid_set(c, set)) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  CoveredDev d = set_dev(c, set);
  if (point_bits) {
    CU(c, cudaMemsetAsync(c->d_tmp_bits, 0, (size_t)c->pts_words * 4, c->stream));
    launch_permute_bits(d.pts, c->d_tmp_bits, c->m.orig, c->m.n, c->stream);
    c->launches += 1;
    CU(c, cudaMemcpyAsync(point_bits, c->d_tmp_bits, (size_t)((c->m.n + 63) / 64) * 8, cudaMemcpyDeviceToHost, c->stream));
  }
  if (voxel_bits)
    CU(c, cudaMemcpyAsync(voxel_bits, d.vox, (size_t)((c->m.n_cov + 63) / 64) * 8, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return ASSCPP_OK;
}

int asscpp_covered_import(AsscppCtx* c, int set, const uint64_t* point_bits, const uint64_t* voxel_bits) {
  if (!c || !valid_set(c, set)) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  CoveredDev d = set_dev(c, set);
  if (point_bits) {
    CU(c, cudaMemcpyAsync(c->d_tmp_bits, point_bits, (size_t)((c->m.n + 63) / 64) * 8, cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaMemsetAsync(d.pts, 0, (size_t)c->pts_words * 4, c->stream));
    launch_permute_bits(c->d_tmp_bits, d.pts, c->m.pos_of, c->m.n, c->stream);
    c->launches += 1;
  }
  if (voxel_bits)
    CU(c, cudaMemcpyAsync(d.vox, voxel_bits, (size_t)((c->m.n_cov + 63) / 64) * 8, cudaMemcpyHostToDevice, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  return ASSCPP_OK;
}

static float percent_of(int64_t matched, int64_t orig) {
  // occlusion_culling.cpp:401,431-432: float counters, float ratio, * 100
  float m = (float)matched, o = (float)orig;
  float ratio = m / o;
  return ratio * 100;
}

int asscpp_coverage_percent(AsscppCtx* c, int set, float* pct) {
  if (!c || !pct) return ASSCPP_EINVAL;
  int64_t nv = 0;
  int rc = asscpp_covered_popcount(c, set, nullptr, &nv);
  if (rc) return rc;
  *pct = percent_of(nv, c->m.n_cov);
  return ASSCPP_OK;
}

int asscpp_coverage_percent_cloud(AsscppCtx* c, const float* xyz, int64_t n, float* pct) {
  if (!c || !pct || (!xyz && n > 0) || n < 0) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  if (n > c->cloud_cap) {
    CU(c, cudaStreamSynchronize(c->stream));
    cudaFree(c->d_cloud);
    c->d_cloud = nullptr; c->cloud_cap = 0;
    int rc = dev_alloc(c, &c->d_cloud, 3 * n, false);
    if (rc) return rc;
    c->cloud_cap = n;
  }
  CoveredDev d = set_dev(c, c->tmp_set);
  CU(c, cudaMemsetAsync(d.vox, 0, (size_t)c->vox_words * 4, c->stream));
  if (n > 0) {
    CU(c, cudaMemcpyAsync(c->d_cloud, xyz, (size_t)n * 12, cudaMemcpyHostToDevice, c->stream));
    launch_cloud_cover(c->d_cloud, n, 1.0f / c->params.cov_leaf, c->cov_min_b, c->cov_div, c->d_cov_keys, c->cov_keys_n,
                       d.vox, c->stream);
    c->launches += 1;
  }
  CU(c, cudaMemsetAsync(c->d_pop, 0, 2 * sizeof(unsigned long long), c->stream));
  launch_popcount(d.vox, c->vox_words, c->d_pop + 1, c->stream);
  c->launches += 1;
  unsigned long long h[2];
  CU(c, cudaMemcpyAsync(h, c->d_pop, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  *pct = percent_of((int64_t)h[1], c->m.n_cov);
  return ASSCPP_OK;
}

int asscpp_covered_or_cloud(AsscppCtx* c, int set, const float* xyz, int64_t n) {
  if (!c || !valid_set(c, set) || (!xyz && n > 0) || n < 0) return ASSCPP_EINVAL;
  if (n == 0) return ASSCPP_OK;
  cudaSetDevice(c->device);
  if (n > c->cloud_cap) {
    CU(c, cudaStreamSynchronize(c->stream));
    cudaFree(c->d_cloud);
    c->d_cloud = nullptr; c->cloud_cap = 0;
    int rc = dev_alloc(c, &c->d_cloud, 3 * n, false);
    if (rc) return rc;
    c->cloud_cap = n;
  }
  CU(c, cudaMemcpyAsync(c->d_cloud, xyz, (size_t)n * 12, cudaMemcpyHostToDevice, c->stream));
  launch_cloud_cover(c->d_cloud, n, 1.0f / c->params.cov_leaf, c->cov_min_b, c->cov_div, c->d_cov_keys, c->cov_keys_n,
                     set_dev(c, set).vox, c->stream);
  c->launches += 1;
  CU(c, cudaGetLastError());
  CU(c, cudaStreamSynchronize(c->stream));  // xyz may be pageable and reused by the caller
  return ASSCPP_OK;
}

int asscpp_covered_invert_voxels(AsscppCtx* c, int set) {
  if (!c || !valid_set(c, set)) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  launch_invert_bits(set_dev(c, set).vox, c->m.n_cov, c->stream);
  c->launches += 1;
  CU(c, cudaGetLastError());
  return ASSCPP_OK;
}

int asscpp_get_visible(AsscppCtx* c, const double pose[7], int32_t* idx_out, int64_t cap, int64_t* n_out) {
  if (!c || !pose || !n_out || cap < 0 || (!idx_out && cap > 0)) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = asscpp_covered_clear(c, c->tmp_set);
  if (rc) return rc;
  rc = asscpp_covered_or_views(c, c->tmp_set, pose, 1, -1);
  if (rc) return rc;
  std::vector<uint64_t> bits((size_t)((c->m.n + 63) / 64) + 1, 0);
  rc = asscpp_covered_export(c, c->tmp_set, bits.data(), nullptr);
  if (rc) return rc;
  rc = fetch_results(c, 1);
  if (rc) return rc;
  int64_t k = 0;
  for (int64_t w = 0; w < (c->m.n + 63) / 64; ++w) {
    uint64_t b = bits[w];
    while (b) {
      int t = __builtin_ctzll(b);
      b &= b - 1;
      if (k < cap) idx_out[k] = (int32_t)(w * 64 + t);
      ++k;
    }
  }
  *n_out = k;
  if (k > cap) { c->err = "idx_out too small"; return ASSCPP_EINVAL; }
  return ASSCPP_OK;
}

// ---- streams / timing --------------------------------------------------------------------------------
int asscpp_sync(AsscppCtx* c) {
  if (!c) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  CU(c, cudaStreamSynchronize(c->stream));
  return ASSCPP_OK;
}

void* asscpp_get_stream(AsscppCtx* c) { return c ? (void*)c->stream : nullptr; }

int asscpp_timer_begin(AsscppCtx* c) {
  if (!c) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  c->timing = true;
  c->kernel_events_used = 0;
  CU(c, cudaEventRecord(c->ev_begin, c->stream));
  return ASSCPP_OK;
}

int asscpp_timer_end(AsscppCtx* c, float* ms) {
  if (!c || !ms) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  CU(c, cudaEventRecord(c->ev_end, c->stream));
  CU(c, cudaEventSynchronize(c->ev_end));
  CU(c, cudaEventElapsedTime(ms, c->ev_begin, c->ev_end));
  float total[3] = {0, 0, 0};
  int64_t calls[3] = {0, 0, 0};
  for (size_t i = 0; i < c->kernel_events_used; ++i) {
    float t = 0;
    CU(c, cudaEventElapsedTime(&t, c->kernel_events[i].e0, c->kernel_events[i].e1));
    total[c->kernel_events[i].kind] += t;
    calls[c->kernel_events[i].kind] += 1;
  }
  c->last_kernel_ms = total[0];
  c->last_kernel_launches = calls[0];
  c->last_coll_ms[0] = total[1]; c->last_coll_ms[1] = total[2];
  c->last_coll_calls[0] = calls[1]; c->last_coll_calls[1] = calls[2];
  c->timing = false;
  c->kernel_events_used = 0;
  return ASSCPP_OK;
}

int64_t asscpp_launch_count(const AsscppCtx* c) { return c ? c->launches : 0; }

int asscpp_debug_phase_cycles(AsscppCtx* c, int enable, uint64_t out[16]) {
  if (!c) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  CU(c, cudaStreamSynchronize(c->stream));
  if (c->d_phase && out) CU(c, cudaMemcpy(out, c->d_phase, 16 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  if (enable && !c->d_phase) {
    int rc = dev_alloc(c, &c->d_phase, 16);
    if (rc) return rc;
  }
  if (c->d_phase) CU(c, cudaMemset(c->d_phase, 0, 16 * sizeof(uint64_t)));
  if (!enable) c->d_phase = nullptr;  // buffer stays in c->allocs and is released with the context
  return ASSCPP_OK;
}

int asscpp_eval_kernel_ms(AsscppCtx* c, float* ms, int64_t* launches) {
  if (!c) return ASSCPP_EINVAL;
  if (ms) *ms = c->last_kernel_ms;
  if (launches) *launches = c->last_kernel_launches;
  return ASSCPP_OK;
}

// ---- uncovered regions ---------------------------------------------------------------------------
static int ensure_region_buffers(AsscppCtx* c) {
  if (c->d_region) return ASSCPP_OK;
  const int64_t n = std::max<int64_t>(c->m.n_cov, 1);
  // [wcount n/32+1 | parent n | root n | size n | box 6n] int32, ids n int32, centres 3n float, total int64
  int rc;
  if ((rc = dev_alloc(c, &c->d_region, (n + 31) / 32 + 1 + 9 * n))) return rc;
  if ((rc = dev_alloc(c, &c->d_region_ids, n))) return rc;
  if ((rc = dev_alloc(c, &c->d_region_ctr, 3 * n))) return rc;
  if ((rc = dev_alloc(c, &c->d_region_total, 1))) return rc;
  if ((rc = dev_alloc(c, &c->d_region_offs, 343))) return rc;
  return ASSCPP_OK;
}

int asscpp_uncovered_voxels(AsscppCtx* c, int set, int32_t* ids, float* centres, int64_t cap, int64_t* n_out) {
  return asscpp_uncovered_voxels_ex(c, set, 0, ids, centres, cap, n_out);
}

int asscpp_uncovered_voxels_ex(AsscppCtx* c, int set, uint32_t flags, int32_t* ids, float* centres, int64_t cap,
                               int64_t* n_out) {
  if (!c || !valid_set(c, set) || cap < 0 || !n_out || set == c->tmp_set) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = ensure_region_buffers(c);
  if (rc) return rc;
  const int64_t n = c->m.n_cov;
  const uint32_t* bits = set_dev(c, set).vox;
  if ((flags & ASSCPP_DIFF_PCL_ALIASING) && n > 0) {
    // quirk Q11: cells the reference's unchecked linear index takes for covered join the set (in the scratch set)
    uint32_t* out = set_dev(c, c->tmp_set).vox;
    launch_covered_alias(bits, out, n, c->d_cov_keys, c->cov_div, (int32_t*)c->d_region_offs, c->stream);
    c->launches += 2;
    bits = out;
  }
  launch_uncovered(bits, n, c->d_region, c->d_region_total, c->d_cov_keys, c->cov_min_b, c->cov_div,
                   c->params.cov_leaf, c->d_region_ids, c->d_region_ctr, n, c->stream);
  c->launches += 3;
  CU(c, cudaGetLastError());
  int64_t total = 0;
  CU(c, cudaMemcpyAsync(&total, c->d_region_total, sizeof(int64_t), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  *n_out = total;
  const int64_t take = std::min(total, cap);
  if (ids && take > 0) CU(c, cudaMemcpy(ids, c->d_region_ids, (size_t)take * 4, cudaMemcpyDeviceToHost));
  if (centres && take > 0) CU(c, cudaMemcpy(centres, c->d_region_ctr, (size_t)take * 12, cudaMemcpyDeviceToHost));
  return ASSCPP_OK;
}

int asscpp_uncovered_clusters(AsscppCtx* c, int set, double tolerance, int32_t min_size, int32_t max_size,
                              int32_t* labels, AsscppCluster* clusters, int64_t cap, int64_t* n_clusters) {
  if (!c || !valid_set(c, set) || !(tolerance > 0) || cap < 0 || !n_clusters || (cap > 0 && !clusters)) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  int rc = ensure_region_buffers(c);
  if (rc) return rc;
  const int64_t n = c->m.n_cov;
  *n_clusters = 0;
  if (n == 0) return ASSCPP_OK;
  // lattice offsets whose centres are closer than the tolerance: <FLANN radiusSearch> keeps squared float32
  // distances strictly below float(tol * tol); centre differences on the lattice are float(d) * leaf
  const float leaf = c->params.cov_leaf, r2 = (float)(tolerance * tolerance);
  const int R = (int)std::floor(tolerance / (double)leaf) + 1;
  if (R > 3) { c->err = "cluster tolerance spans more than 3 coverage voxels"; return ASSCPP_ERANGE; }
  std::vector<int3> offs;
  for (int dz = -R; dz <= R; ++dz)
    for (int dy = -R; dy <= R; ++dy)
      for (int dx = -R; dx <= R; ++dx) {
        if (!dx && !dy && !dz) continue;
        const float fx = (float)dx * leaf, fy = (float)dy * leaf, fz = (float)dz * leaf;
        const float d2 = fx * fx + fy * fy + fz * fz;
        if (d2 < r2) offs.push_back(make_int3(dx, dy, dz));
      }
  int32_t* wcount = c->d_region;
  int32_t* parent = wcount + (n + 31) / 32 + 1;
  int32_t *root = parent + n, *size = root + n, *box = size + n;
  if (!offs.empty())
    CU(c, cudaMemcpyAsync(c->d_region_offs, offs.data(), offs.size() * sizeof(int3), cudaMemcpyHostToDevice, c->stream));
  launch_clusters(set_dev(c, set).vox, n, c->d_cov_keys, c->cov_min_b, c->cov_div, c->d_region_offs, (int)offs.size(),
                  parent, root, size, box, c->stream);
  c->launches += 4;
  CU(c, cudaGetLastError());
  std::vector<int32_t> h_root((size_t)n), h_size((size_t)n), h_box((size_t)6 * n);
  CU(c, cudaMemcpyAsync(h_root.data(), root, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemcpyAsync(h_size.data(), size, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaMemcpyAsync(h_box.data(), box, (size_t)n * 24, cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  // clusters = roots whose size is in range (heuristic.cpp:157-158), largest first (<PCL> sorts by size)
  std::vector<int32_t> roots;
  for (int64_t i = 0; i < n; ++i)
    if (h_root[(size_t)i] == (int32_t)i && h_size[(size_t)i] >= min_size && h_size[(size_t)i] <= max_size) roots.push_back((int32_t)i);
  std::stable_sort(roots.begin(), roots.end(), [&](int32_t a, int32_t b) { return h_size[(size_t)a] > h_size[(size_t)b]; });
  *n_clusters = (int64_t)roots.size();
  std::vector<int32_t> rank((size_t)n, -1);
  for (size_t k = 0; k < roots.size(); ++k) rank[(size_t)roots[k]] = (int32_t)k;
  if (labels)
    for (int64_t i = 0; i < n; ++i) labels[i] = h_root[(size_t)i] >= 0 ? rank[(size_t)h_root[(size_t)i]] : -1;
  const int64_t take = std::min<int64_t>(cap, (int64_t)roots.size());
  for (int64_t k = 0; k < take; ++k) {
    AsscppCluster& o = clusters[k];
    const int32_t r = roots[(size_t)k];
    o.size = h_size[(size_t)r];
    o.first_id = r;  // roots are the smallest id of their component (hooking is always larger -> smaller)
    float cmin[3], cmax[3];
    for (int a = 0; a < 3; ++a) {
      o.min_ijk[a] = h_box[(size_t)6 * r + a];
      o.max_ijk[a] = h_box[(size_t)6 * r + 3 + a];
      // findClusterBB :224-233: the cluster's own voxel grid; centres of its min and max cells
      const float b_min = (float)o.min_ijk[a] * leaf;
      cmin[a] = (b_min + leaf * 0.5f) + (float)0 * leaf;
      cmax[a] = (b_min + leaf * 0.5f) + (float)(o.max_ijk[a] - o.min_ijk[a]) * leaf;
    }
    const double maximizeSizeXY = 5, maximizeSizeZ = 3;  // :236-237
    o.grid_size[0] = std::abs(cmax[0] - cmin[0]) + maximizeSizeXY;
    o.grid_size[1] = std::abs(cmax[1] - cmin[1]) + maximizeSizeXY;
    o.grid_size[2] = std::abs(cmax[2] - cmin[2]) + maximizeSizeZ;
    o.grid_start[0] = cmin[0] - double(maximizeSizeXY / 2);
    o.grid_start[1] = cmin[1] - double(maximizeSizeXY / 2);
    o.grid_start[2] = cmin[2];
  }
  return ASSCPP_OK;
}

// ---- multi-GPU -----------------------------------------------------------------------------------------
int asscpp_comm_unique_id(uint8_t id_out[128]) {
  std::string err;
  if (!id_out) return ASSCPP_EINVAL;
  if (!g_nccl.load(&err)) { g_create_error = err; return ASSCPP_ENCCL; }
  NcclUniqueId id;
  int r = g_nccl.GetUniqueId(&id);
  if (r != 0) { g_create_error = "ncclGetUniqueId failed"; return ASSCPP_ENCCL; }
  memcpy(id_out, id.internal, 128);
  return ASSCPP_OK;
}

int asscpp_comm_init(AsscppCtx* c, int nranks, int rank, const uint8_t id[128]) {
  if (!c || !id || nranks < 1 || rank < 0 || rank >= nranks) return ASSCPP_EINVAL;
  cudaSetDevice(c->device);
  if (!g_nccl.load(&c->err)) return ASSCPP_ENCCL;
  NcclUniqueId uid;
  memcpy(uid.internal, id, 128);
  int r = g_nccl.CommInitRank(&c->comm, nranks, uid, rank);
  if (r != 0) {
    c->err = std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "error");
    c->comm = nullptr;
    return ASSCPP_ENCCL;
  }
  c->nranks = nranks;
  c->rank = rank;
  return ASSCPP_OK;
}

// one process, several GPUs (asscpp_comm_init(ctxs[], n)): ncclCommInitAll over the contexts' devices
int asscpp_comm_init_all(AsscppCtx** ctxs, int n) {
  if (!ctxs || n < 1) return ASSCPP_EINVAL;
  for (int i = 0; i < n; ++i)
    if (!ctxs[i] || ctxs[i]->comm) return ASSCPP_EINVAL;
  if (n == 1) { ctxs[0]->nranks = 1; ctxs[0]->rank = 0; return ASSCPP_OK; }
  if (!g_nccl.load(&ctxs[0]->err)) return ASSCPP_ENCCL;
  if (!g_nccl.CommInitAll) { ctxs[0]->err = "libnccl lacks ncclCommInitAll"; return ASSCPP_ENCCL; }
  std::vector<NcclComm> comms((size_t)n, nullptr);
  std::vector<int> devs((size_t)n);
  for (int i = 0; i < n; ++i) devs[(size_t)i] = ctxs[i]->device;
  int r = g_nccl.CommInitAll(comms.data(), n, devs.data());
  if (r != 0) {
    ctxs[0]->err = std::string("ncclCommInitAll: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "error");
    return ASSCPP_ENCCL;
  }
  for (int i = 0; i < n; ++i) { ctxs[i]->comm = comms[(size_t)i]; ctxs[i]->nranks = n; ctxs[i]->rank = i; }
  return ASSCPP_OK;
}

static int ensure_gather(AsscppCtx* c, size_t bytes) {
  if (bytes <= c->gather_bytes) return ASSCPP_OK;
  CU(c, cudaStreamSynchronize(c->stream));
  cudaFree(c->d_gather);
  c->d_gather = nullptr; c->gather_bytes = 0;
  uint8_t* p;
  int rc = dev_alloc(c, &p, (int64_t)bytes, false);
  if (rc) return rc;
  c->d_gather = p;
  c->gather_bytes = bytes;
  return ASSCPP_OK;
}

// ncclAllGather needs the same count on every rank: exchange the counts first (8 bytes per rank) and refuse a
// mismatch instead of hanging or corrupting the gather buffer.  Costs one small collective and a stream sync, so it
// is done by the host-output variants only (they synchronise anyway).
static int check_equal_counts(AsscppCtx* c, int64_t nviews_local) {
  if (!c->d_cnt) {
    int rc = dev_alloc(c, &c->d_cnt, 1 + c->nranks, false);
    if (rc) return rc;
  }
  std::vector<int64_t> h((size_t)c->nranks + 1, 0);
  h[0] = nviews_local;
  CU(c, cudaMemcpyAsync(c->d_cnt, h.data(), sizeof(int64_t), cudaMemcpyHostToDevice, c->stream));
  int r = g_nccl.AllGather(c->d_cnt, c->d_cnt + 1, sizeof(int64_t), kNcclUint8, c->comm, c->stream);
  if (r != 0) { c->err = "ncclAllGather failed"; return ASSCPP_ENCCL; }
  CU(c, cudaMemcpyAsync(h.data() + 1, c->d_cnt + 1, (size_t)c->nranks * sizeof(int64_t), cudaMemcpyDeviceToHost, c->stream));
  CU(c, cudaStreamSynchronize(c->stream));
  for (int k = 0; k < c->nranks; ++k)
    if (h[(size_t)k + 1] != nviews_local) {
      c->err = "allgather needs the same view count on every rank (pad the shards): rank " + std::to_string(k) +
               " has " + std::to_string(h[(size_t)k + 1]) + ", this rank " + std::to_string(nviews_local);
      return ASSCPP_EINVAL;
    }
  return ASSCPP_OK;
}

int asscpp_allgather_results(AsscppCtx* c, int64_t nviews_local, AsscppViewResult* out) {
  if (!c || nviews_local < 0) return ASSCPP_EINVAL;
  if (!c->comm) { c->err = "communicator not initialised"; return ASSCPP_ESTATE; }
  if (nviews_local > c->last_nviews) { c->err = "more results requested than views evaluated"; return ASSCPP_EINVAL; }
  cudaSetDevice(c->device);
  if (out) {
    int rc = check_equal_counts(c, nviews_local);
    if (rc) return rc;
  }
  size_t bytes = (size_t)nviews_local * sizeof(AsscppViewResult);
  int rc = ensure_gather(c, bytes * c->nranks);
  if (rc) return rc;
  if (bytes) {
    AsscppCtx::TimedSpan* sp = span_begin(c, 1);
    int r = g_nccl.AllGather(c->d_results, c->d_gather, bytes, kNcclUint8, c->comm, c->stream);
    span_end(c, sp);
    if (r != 0) { c->err = "ncclAllGather failed"; return ASSCPP_ENCCL; }
    if (out) CU(c, cudaMemcpyAsync(out, c->d_gather, bytes * c->nranks, cudaMemcpyDeviceToHost, c->stream));
  }
  if (out) CU(c, cudaStreamSynchronize(c->stream));  // out == NULL: gathered scores stay on the device, async
  return ASSCPP_OK;
}

int asscpp_allgather_gain(AsscppCtx* c, int64_t nviews_local, uint32_t* gain_pts, uint32_t* gain_vox) {
  if (!c || nviews_local < 0) return ASSCPP_EINVAL;
  if (!c->comm) { c->err = "communicator not initialised"; return ASSCPP_ESTATE; }
  if (nviews_local > c->last_nviews) { c->err = "more results requested than views evaluated"; return ASSCPP_EINVAL; }
  cudaSetDevice(c->device);
  const bool to_host = gain_pts || gain_vox;
  if (to_host) {
    int rc = check_equal_counts(c, nviews_local);
    if (rc) return rc;
  }
  const size_t bytes = (size_t)nviews_local * sizeof(uint2);
  int rc = ensure_gather(c, std::max<size_t>(bytes * c->nranks, 16));
  if (rc) return rc;
  if (bytes) {
    AsscppCtx::TimedSpan* sp = span_begin(c, 1);
    int r = g_nccl.AllGather(c->d_gains, c->d_gather, bytes, kNcclUint8, c->comm, c->stream);
    span_end(c, sp);
    if (r != 0) { c->err = "ncclAllGather failed"; return ASSCPP_ENCCL; }
  }
  if (!to_host) return ASSCPP_OK;  // gathered scores stay on the device (rank-major uint2 array), async
  const int64_t total = nviews_local * c->nranks;
  rc = fetch_gains(c, (const uint2*)c->d_gather, total);
  if (rc) return rc;
  for (int64_t v = 0; v < total; ++v) {
    if (gain_pts) gain_pts[v] = c->h_gains[(size_t)v].x;
    if (gain_vox) gain_vox[v] = c->h_gains[(size_t)v].y;
  }
  return ASSCPP_OK;
}

int asscpp_allreduce_or(AsscppCtx* c, int set) {
  if (!c || !valid_set(c, set)) return ASSCPP_EINVAL;
  if (!c->comm) { c->err = "communicator not initialised"; return ASSCPP_ESTATE; }
  cudaSetDevice(c->device);
  size_t bytes = (size_t)set_words(c) * 4;
  int rc = ensure_gather(c, bytes * c->nranks);
  if (rc) return rc;
  // NCCL has no bitwise-OR reduction: gather every rank's bitset, then OR them locally
  AsscppCtx::TimedSpan* sp = span_begin(c, 2);
  int r = g_nccl.AllGather(c->sets[set].buf, c->d_gather, bytes, kNcclUint8, c->comm, c->stream);
  if (r != 0) { c->err = "ncclAllGather failed"; return ASSCPP_ENCCL; }
  launch_or_reduce(c->sets[set].buf, (const uint32_t*)c->d_gather, set_words(c), c->nranks, c->stream);
  span_end(c, sp);
  c->launches += 1;
  CU(c, cudaGetLastError());
  return ASSCPP_OK;
}

int asscpp_collective_ms(AsscppCtx* c, float* allgather_ms, int64_t* allgather_calls, float* allreduce_or_ms,
                         int64_t* allreduce_or_calls) {
  if (!c) return ASSCPP_EINVAL;
  if (allgather_ms) *allgather_ms = c->last_coll_ms[0];
  if (allgather_calls) *allgather_calls = c->last_coll_calls[0];
  if (allreduce_or_ms) *allreduce_or_ms = c->last_coll_ms[1];
  if (allreduce_or_calls) *allreduce_or_calls = c->last_coll_calls[1];
  return ASSCPP_OK;
}

int asscpp_comm_destroy(AsscppCtx* c) {
  if (!c) return ASSCPP_EINVAL;
  if (c->comm && g_nccl.CommDestroy) {
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    g_nccl.CommDestroy(c->comm);
  }
  c->comm = nullptr;
  c->nranks = 1;
  c->rank = 0;
  cudaFree(c->d_cnt);
  c->d_cnt = nullptr;
  return ASSCPP_OK;
}

int asscpp_frustum_planes(const AsscppParams* p, const double pose[7], float planes[24], float origin[3]) {
  if (!p || !pose || !planes || !origin) return ASSCPP_EINVAL;
  frustum_planes_host(*p, pose, planes, origin, nullptr);
  return ASSCPP_OK;
}

int asscpp_pose_rpy(const double pose[7], double rpy[3]) {
  if (!pose || !rpy) return ASSCPP_EINVAL;
  pose_rpy_host(pose, rpy);
  return ASSCPP_OK;
}

int asscpp_viewport_transform(const double pose[7], const float* xyz, int64_t n, float* out) {
  // OcclusionCulling::pointCloudViewportTransform (occlusion_culling.cpp:538-618): translate by -camera position,
  // rotate about z by -yaw and about y by -pitch; every stage in double on float32 inputs, narrowed to float32
  // (transformPointMatVec, :521-531).  Same arithmetic as acc_point in kernels.cu.
  if (!pose || (!xyz && n > 0) || (!out && n > 0) || n < 0) return ASSCPP_EINVAL;
  ViewTrig t;
  viewport_trig_host(pose, &t);
  for (int64_t i = 0; i < n; ++i) {
    const float a0 = (float)((double)xyz[3 * i] + t.t[0]), a1 = (float)((double)xyz[3 * i + 1] + t.t[1]),
                a2 = (float)((double)xyz[3 * i + 2] + t.t[2]);
    const float b0 = (float)(t.cz * (double)a0 + (-t.sz) * (double)a1), b1 = (float)(t.sz * (double)a0 + t.cz * (double)a1);
    out[3 * i] = (float)(t.cy * (double)b0 + t.sy * (double)a2);
    out[3 * i + 1] = b1;
    out[3 * i + 2] = (float)((-t.sy) * (double)b0 + t.cy * (double)a2);
  }
  return ASSCPP_OK;
}

}  // extern "C"
