// multi.cpp -- one PROCESS, several GPUs: the reference is a single process (cscpp/src/coverage_heuristic_test.cpp:42),
// so a caller that links the C++ shim cannot start one rank per GPU.  AsscppMulti holds one context per device
// (model replicated), one host worker thread per device, and shards every batch of viewpoints over them:
// contiguous blocks of whole groups, no data-path collective; the covered-set exchange is the same NCCL OR-allreduce
// the multi-process path uses (asscpp_allreduce_or), set up with ncclCommInitAll (asscpp_comm_init_all).
// Host code only: everything goes through the public C ABI.
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/asscpp_gpu.h"

struct AsscppMulti {
  std::vector<AsscppCtx*> ctx;
  std::vector<std::thread> workers;
  std::vector<std::function<int()> > task;
  std::vector<int> rc;
  std::vector<int> state;  // 0 idle, 1 task posted, 2 done
  std::mutex mu;
  std::condition_variable cv_post, cv_done;
  bool stop = false;
  std::string err;
};

namespace {

void worker(AsscppMulti* m, int i) {
  for (;;) {
    std::function<int()> fn;
    {
      std::unique_lock<std::mutex> lk(m->mu);
      m->cv_post.wait(lk, [&] { return m->stop || m->state[(size_t)i] == 1; });
      if (m->stop) return;
      fn = m->task[(size_t)i];
    }
    const int rc = fn();
    {
      std::lock_guard<std::mutex> lk(m->mu);
      m->rc[(size_t)i] = rc;
      m->state[(size_t)i] = 2;
    }
    m->cv_done.notify_all();
  }
}

// run f(i) on every device's worker thread, wait for all; first non-zero code wins
int run_all(AsscppMulti* m, const std::function<int(int)>& f) {
  const int n = (int)m->ctx.size();
  {
    std::lock_guard<std::mutex> lk(m->mu);
    for (int i = 0; i < n; ++i) {
      m->task[(size_t)i] = [f, i] { return f(i); };
      m->state[(size_t)i] = 1;
    }
  }
  m->cv_post.notify_all();
  std::unique_lock<std::mutex> lk(m->mu);
  m->cv_done.wait(lk, [&] {
    for (int i = 0; i < n; ++i)
      if (m->state[(size_t)i] != 2) return false;
    return true;
  });
  int rc = 0;
  for (int i = 0; i < n; ++i) {
    m->state[(size_t)i] = 0;
    if (m->rc[(size_t)i] && !rc) {
      rc = m->rc[(size_t)i];
      m->err = std::string("device ") + std::to_string(i) + ": " + asscpp_last_error(m->ctx[(size_t)i]);
    }
  }
  return rc;
}

// views [lo, hi) of device i: contiguous blocks of whole groups, the first (ngroups mod ndev) devices get one more
void shard(int64_t nviews, int group, int ndev, int i, int64_t* lo, int64_t* hi) {
  const int64_t ngroups = (nviews + group - 1) / group;
  const int64_t base = ngroups / ndev, extra = ngroups % ndev;
  const int64_t g0 = i * base + (i < extra ? i : extra), g1 = g0 + base + (i < extra ? 1 : 0);
  *lo = g0 * group < nviews ? g0 * group : nviews;
  *hi = g1 * group < nviews ? g1 * group : nviews;
}

}  // namespace

extern "C" {

int asscpp_multi_create(const int* devices, int ndev, const float* xyz, int64_t n, const AsscppParams* params,
                        AsscppMulti** out) {
  if (!out || !devices || ndev < 1 || !params) return ASSCPP_EINVAL;
  *out = nullptr;
  AsscppMulti* m = new AsscppMulti();
  m->ctx.assign((size_t)ndev, nullptr);
  // the contexts are built concurrently (the host-side layout of the model is the expensive part)
  std::vector<int> rcs((size_t)ndev, 0);
  std::vector<std::thread> th;
  for (int i = 0; i < ndev; ++i)
    th.push_back(std::thread([&, i] { rcs[(size_t)i] = asscpp_create(devices[i], xyz, n, params, &m->ctx[(size_t)i]); }));
  for (size_t i = 0; i < th.size(); ++i) th[i].join();
  int rc = 0;
  for (int i = 0; i < ndev; ++i)
    if (rcs[(size_t)i] && !rc) rc = rcs[(size_t)i];
  if (!rc && ndev > 1) rc = asscpp_comm_init_all(m->ctx.data(), ndev);
  if (rc) {
    for (int i = 0; i < ndev; ++i) asscpp_destroy(m->ctx[(size_t)i]);
    delete m;
    return rc;
  }
  m->task.resize((size_t)ndev);
  m->rc.assign((size_t)ndev, 0);
  m->state.assign((size_t)ndev, 0);
  for (int i = 0; i < ndev; ++i) m->workers.push_back(std::thread(worker, m, i));
  *out = m;
  return ASSCPP_OK;
}

int asscpp_multi_destroy(AsscppMulti* m) {
  if (!m) return ASSCPP_OK;
  {
    std::lock_guard<std::mutex> lk(m->mu);
    m->stop = true;
  }
  m->cv_post.notify_all();
  for (size_t i = 0; i < m->workers.size(); ++i) m->workers[i].join();
  for (size_t i = 0; i < m->ctx.size(); ++i) {
    asscpp_comm_destroy(m->ctx[i]);
    asscpp_destroy(m->ctx[i]);
  }
  delete m;
  return ASSCPP_OK;
}

int asscpp_multi_num_devices(const AsscppMulti* m) { return m ? (int)m->ctx.size() : 0; }
AsscppCtx* asscpp_multi_context(AsscppMulti* m, int i) {
  return (m && i >= 0 && i < (int)m->ctx.size()) ? m->ctx[(size_t)i] : nullptr;
}
const char* asscpp_multi_last_error(const AsscppMulti* m) { return m ? m->err.c_str() : "null AsscppMulti"; }

// covered sets are mirrored: the same handle names the same set on every device
int asscpp_multi_covered_new(AsscppMulti* m, int* set_out) {
  if (!m || !set_out) return ASSCPP_EINVAL;
  std::vector<int> ids(m->ctx.size(), -1);
  int rc = run_all(m, [&](int i) { return asscpp_covered_new(m->ctx[(size_t)i], &ids[(size_t)i]); });
  if (rc) return rc;
  for (size_t i = 1; i < ids.size(); ++i)
    if (ids[i] != ids[0]) { m->err = "covered-set handles diverged between devices (use the asscpp_multi_* calls only)"; return ASSCPP_ESTATE; }
  *set_out = ids[0];
  return ASSCPP_OK;
}
int asscpp_multi_covered_free(AsscppMulti* m, int set) {
  if (!m) return ASSCPP_EINVAL;
  return run_all(m, [&](int i) { return asscpp_covered_free(m->ctx[(size_t)i], set); });
}
int asscpp_multi_covered_clear(AsscppMulti* m, int set) {
  if (!m) return ASSCPP_EINVAL;
  return run_all(m, [&](int i) { return asscpp_covered_clear(m->ctx[(size_t)i], set); });
}
int asscpp_multi_allreduce_or(AsscppMulti* m, int set) {
  if (!m) return ASSCPP_EINVAL;
  if (m->ctx.size() == 1) return ASSCPP_OK;
  return run_all(m, [&](int i) {
    int rc = asscpp_allreduce_or(m->ctx[(size_t)i], set);
    return rc ? rc : asscpp_sync(m->ctx[(size_t)i]);
  });
}
int asscpp_multi_covered_popcount(AsscppMulti* m, int set, int64_t* n_points, int64_t* n_voxels) {
  if (!m) return ASSCPP_EINVAL;
  return asscpp_covered_popcount(m->ctx[0], set, n_points, n_voxels);  // every device holds the same set
}
int asscpp_multi_coverage_percent(AsscppMulti* m, int set, float* pct) {
  if (!m) return ASSCPP_EINVAL;
  return asscpp_coverage_percent(m->ctx[0], set, pct);
}

// asscpp_eval_ex over all devices.  results / gain_pts / gain_vox (any may be NULL) receive nviews entries in the order
// of poses.  With opts->accum_set >= 0 the accumulator is OR-allreduced afterwards: every device ends with the union.
int asscpp_multi_eval(AsscppMulti* m, const double* poses, int64_t nviews, const AsscppEvalOpts* opts,
                      AsscppViewResult* results, uint32_t* gain_pts, uint32_t* gain_vox) {
  if (!m || !opts || (!poses && nviews > 0) || nviews < 0 || opts->poses_on_device) return ASSCPP_EINVAL;
  const int ndev = (int)m->ctx.size();
  const int group = opts->group_size > 0 ? opts->group_size : 1;
  int rc = run_all(m, [&](int i) {
    int64_t lo, hi;
    shard(nviews, group, ndev, i, &lo, &hi);
    AsscppCtx* c = m->ctx[(size_t)i];
    int r = asscpp_eval_ex(c, poses + 7 * lo, hi - lo, opts);
    if (r) return r;
    if (results && (r = asscpp_get_results(c, results + lo, hi - lo))) return r;
    if ((gain_pts || gain_vox) && (r = asscpp_get_gain(c, gain_pts ? gain_pts + lo : nullptr, gain_vox ? gain_vox + lo : nullptr, hi - lo))) return r;
    if (opts->accum_set >= 0 && ndev > 1 && (r = asscpp_allreduce_or(c, opts->accum_set))) return r;
    return asscpp_sync(c);
  });
  return rc;
}

// asscpp_covered_or_views over all devices + OR-allreduce
int asscpp_multi_covered_or_views(AsscppMulti* m, int set, const double* poses, int64_t nviews, int64_t min_visible) {
  if (!m || (!poses && nviews > 0) || nviews < 0) return ASSCPP_EINVAL;
  const int ndev = (int)m->ctx.size();
  return run_all(m, [&](int i) {
    int64_t lo, hi;
    shard(nviews, 1, ndev, i, &lo, &hi);
    AsscppCtx* c = m->ctx[(size_t)i];
    int r = asscpp_covered_or_views(c, set, poses + 7 * lo, hi - lo, min_visible);
    if (r) return r;
    if (ndev > 1 && (r = asscpp_allreduce_or(c, set))) return r;
    return asscpp_sync(c);
  });
}

}  // extern "C"
