#!/usr/bin/env python
"""bench.py -- viewpoint-coverage evaluation throughput (BASELINE.json metric) on N B200s.

  python bench.py --gpus N --steps K --warmup W [--config 2|3|5]   # this repo's CUDA path
  python bench.py --impl reference --steps K --warmup W            # the reference algorithm on the host cores

Workloads (config.workload), BASELINE.json `configs`:
  --config 3 (default, the configuration the metric is quoted on): synthetic 10M-point structure
             (synth_structure_v1), 100k candidate viewpoints per GPU (synth_views_v1), 0.05 m occlusion grid;
  --config 2: 1M points, 10k viewpoints, 0.5 m grid;   --config 5: 50M points, 1M viewpoints, 0.05 m grid.

A "step" is one pass of the hot path over the configuration's whole viewpoint batch, the way the search loops use
it (sspp/src/astar.cpp:311-357, sspp/src/pathplanner.cpp:767-848):
  1. every viewpoint of the batch: frustum cull -> per-view bit-grid -> 3D-DDA -> gain against the RUNNING covered
     set (popcount of visible & ~covered, points and coverage voxels);
  2. the scores (8 bytes per view) come back to the host -- gathered over the ranks with NCCL when N > 1;
  3. the best-scoring viewpoint is accepted: its visible set is OR-ed into the covered set on the rank that owns it;
  4. N > 1: covered-bitset OR-allreduce (NCCL allgather + fused OR kernel), so every rank scores the next step
     against the same set.
N > 1: viewpoints are sharded over the ranks (weak scaling: every rank gets its own batch), the point SoA and the
covered set are replicated.

`value`  : VE/s with poses resident in HBM, device-timed (CUDA events on the launching stream), max over ranks.
`e2e`    : same metric through the host-buffer C-ABI call (asscpp_eval_ex + asscpp_get_gain / asscpp_allgather_gain),
           H2D of the poses from pinned memory and D2H of the scores inside the timed region, wall clock.
`roofline`: for the dominant kernel (eval_views_kernel).  `frac` follows the contract: algorithmic bytes per VE
           (B_VE = 16 N + 20 S_v + G_v/8 + 2 N/8 + 36) / launch time / measured HBM peak.  The kernel
           pre-culls voxel buckets and never reads most of the cloud, so frac > 1 means bytes AVOIDED; beside it:
           `dram` (bytes the kernel really moved, from the ncu capture of the same workload under profiles/),
           `fp32` (the reference algorithm's 48 flop per point-view test against the FP32 peak) and `issue`
           (warp instructions per launch from the same capture / this run's launch time / SMs x sampled clock).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "viewpoint_coverage_evals_per_sec"
UNIT = "VE/s"
SEED_STRUCTURE, SEED_POINTS, SEED_VIEWS = 0xA55C0001, 0xA55C0002, 0xA55C0003

CONFIGS = {
    # points, views per step per GPU, leaf, default timed steps (>= 2 s of GPU work), BASELINE.json index
    2: dict(points=1_000_000, views=10_000, leaf=0.5, steps=300, name="configs[1]: 10k viewpoints vs 1M points, 0.5 m grid"),
    3: dict(points=10_000_000, views=100_000, leaf=0.05, steps=20,
            name="configs[2] per GPU: 100k viewpoints vs 10M points, 0.05 m grid"),
    5: dict(points=50_000_000, views=1_000_000, leaf=0.05, steps=2,
            name="configs[4] per GPU: 1M-viewpoint sweep vs 50M points, 0.05 m grid"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", type=int, default=3, choices=sorted(CONFIGS))
    ap.add_argument("--points", type=int, default=None)
    ap.add_argument("--views-per-step", type=int, default=None)
    ap.add_argument("--leaf", type=float, default=None)
    ap.add_argument("--cov-leaf", type=float, default=0.5)
    ap.add_argument("--brute", action="store_true", help="disable the bucket pre-cull (stream all points per view)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of each cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--port", action="store_true", help="reference arm: time the oracle port even when oracle/_ref exists")
    a = ap.parse_args()
    c = CONFIGS[a.config]
    a.points = a.points or c["points"]
    a.views_per_step = a.views_per_step or c["views"]
    a.leaf = a.leaf or c["leaf"]
    a.steps = a.steps or c["steps"]
    a.warmup = max(a.warmup, 3)
    return a


def workload_config(a):
    """Identical for both arms (the driver compares it); per-arm details go under `arm` on each line."""
    return {
        "workload": "synth_structure_v1 %d points x %d viewpoints per step per GPU, leaf %.3g m, cov leaf %.3g m, "
                    "frustum 58x45 deg 0.7-6 m; step = score all viewpoints against the running covered set, "
                    "accept the best, update the set" % (a.points, a.views_per_step, a.leaf, a.cov_leaf),
        "baseline_config": CONFIGS[a.config]["name"],
        "points": a.points, "views_per_step_per_gpu": a.views_per_step, "leaf_m": a.leaf, "cov_leaf_m": a.cov_leaf,
        "covered_set": "running (accepted viewpoints of the warm-up and of every earlier step)",
    }


def workload_key(a):
    return "%dpts_%dviews_leaf%g_%s" % (a.points, a.views_per_step, a.leaf, "brute" if a.brute else "bucketed")


# ---- host-only input generators (asscpp_b200/lib/libasscpp_host.so: no CUDA inside) -------------------------------
def host_lib():
    p = os.path.join(ROOT, "asscpp_b200", "lib", "libasscpp_host.so")
    if not os.path.exists(p):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "asscpp_b200", "csrc"), "../lib/libasscpp_host.so"])
    L = ctypes.CDLL(p)
    L.asscpp_synth_structure_v1.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_void_p]
    L.asscpp_synth_views_v1.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_void_p]
    return L


def make_inputs(a, rank):
    L = host_lib()
    xyz = np.empty((a.points, 3), np.float32)
    if L.asscpp_synth_structure_v1(SEED_STRUCTURE, SEED_POINTS, a.points, xyz.ctypes.data):
        raise SystemExit("synth_structure failed")
    pool = np.empty((a.views_per_step, 7), np.float64)
    # every rank draws its own viewpoints (rank 0 = the canonical synth_views_v1 stream)
    if L.asscpp_synth_views_v1(SEED_STRUCTURE, SEED_VIEWS + rank, a.views_per_step, pool.ctypes.data):
        raise SystemExit("synth_views failed")
    return xyz, pool


def step_offset(s, nviews):
    """Step s evaluates the rank's batch rotated by this offset (same viewpoints, another order)."""
    return (s * 7919) % max(1, nviews)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            t = [x.strip() for x in r.split(",")]
            if len(t) < 7:
                continue
            try:
                sm.append(float(t[0]))
                mx.append(float(t[1]))
            except ValueError:
                continue
            for nm, val in zip(names, t[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_capture(a):
    """ncu figures of eval_views_kernel for THIS workload (keyed by size and cull mode), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(p)).get("captures", {}).get(workload_key(a))
    except Exception:
        return None


def algorithmic_bytes(n_points, res):
    """B_VE = 16 N + 20 S_v + G_v/8 + 2 N/8 + 36 summed over the views."""
    v = len(res)
    return (16.0 * n_points + 2.0 * n_points / 8.0 + 36.0) * v + 20.0 * float(res["survivors"].sum(dtype=np.float64)) \
        + float(res["cells"].sum(dtype=np.float64)) / 8.0


def best_view(gp, gv):
    """Greedy acceptance: most newly covered coverage voxels, then most newly covered points, then lowest index."""
    key = gv.astype(np.int64) * (1 << 32) + gp.astype(np.int64)
    return int(np.argmax(key))


# ---- CPU legs (oracle = CPU restatement of the reference algorithm; test infrastructure, used here as the baseline) --
def cpu_sample(orc, O, poses, seconds, nthreads, cov):
    probe = poses[:max(nthreads, 2)]
    t = time.perf_counter()
    orc.eval_batch(probe, memo=True, nthreads=nthreads, covered_pts=cov[0], covered_vox=cov[1])
    per_round = max(time.perf_counter() - t, 1e-4)
    k = int(max(len(probe), min(len(poses), len(probe) * max(1, round(seconds / per_round)))))
    sample = poses[:k]
    t = time.perf_counter()
    st = orc.eval_batch(sample, memo=True, nthreads=nthreads, covered_pts=cov[0], covered_vox=cov[1])
    dt = time.perf_counter() - t
    return {"views": k, "seconds": dt, "ve_per_s": k / dt, "point_rays_per_s": float(st[:, O.ST_SURVIVORS].sum()) / dt}


def run_reference(a):
    """The reference arm: the reference's CPU implementation of the path on all host threads.
    kind "reference": oracle/_ref/libasscpp_ref.so -- the reference's OWN OcclusionCulling / FrustumCullingTT /
    VoxelGridOcclusionEstimationT sources compiled from /root/reference against the stand-in third-party headers
    (oracle/Makefile), one OcclusionCulling object per host thread (the class is not re-entrant).  Where that library
    does not exist, or the model is too large to hold one copy per thread, kind "port": the oracle restatement of the
    same algorithm.  Loads oracle/ and the host-only generator library, never the product's CUDA library."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from oracle import oracle as O
    from oracle import ref as R
    nthreads = os.cpu_count() or 1
    xyz, pool = make_inputs(a, 0)
    orc = O.Oracle(xyz, O.make_params(leaf=a.leaf, cov_leaf=a.cov_leaf))
    cov_p = np.zeros(orc.pts_words(), np.uint64)
    cov_v = np.zeros(orc.vox_words(), np.uint64)
    use_ref = os.path.exists(R.LIB_PATH) and a.points * 40 * nthreads < 0.5 * host_memory_bytes() and not a.port
    kind = "reference" if use_ref else "port"
    if use_ref:
        roc = R.RefOcclusionCulling(xyz=xyz, leaf=a.leaf, nthreads=nthreads)
        pt_cov = orc.cov_ids()

        def evaluate(batch):
            st = roc.eval_batch_gain(batch, cov_p, cov_v, pt_cov, orc.n_cov)
            return st[:, 0], st[:, 3], st[:, 4]

        def accept(pose):
            roc.accumulate(pose, cov_p, cov_v, pt_cov)
    else:
        def evaluate(batch):
            st = orc.eval_batch(batch, memo=True, nthreads=nthreads, covered_pts=cov_p, covered_vox=cov_v)
            return st[:, O.ST_SURVIVORS], st[:, O.ST_GAIN_PTS], st[:, O.ST_GAIN_VOX]

        def accept(pose):
            orc.accumulate(pose.reshape(1, 7), cov_p, cov_v, min_visible=-1, nthreads=1)
    # bounded sample per step: the whole run stays within a few minutes
    t = time.perf_counter()
    evaluate(pool[:nthreads])
    per_round = max(time.perf_counter() - t, 1e-4)
    total_steps = a.steps + a.warmup
    rounds = max(1, int(round(max(a.cpu_seconds, 10.0) * 5 / total_steps / per_round)))
    k = min(nthreads * rounds, a.views_per_step)
    times, rays = [], 0.0
    for s in range(total_steps):
        lo = step_offset(s, a.views_per_step)
        batch = np.ascontiguousarray(np.roll(pool, -lo, axis=0)[:k])
        t = time.perf_counter()
        surv, gp, gv = evaluate(batch)
        w = best_view(gp, gv)
        accept(batch[w])
        dt = time.perf_counter() - t
        if s >= a.warmup:
            times.append(dt)
            rays += float(surv.sum())
    tot = sum(times)
    value = k * a.steps / tot
    what = ("oracle/_ref/libasscpp_ref.so: the reference's own cscpp sources (compiled against stand-in PCL/Eigen/tf "
            "headers), one OcclusionCulling object per thread, gain as a bitset popcount" if use_ref else
            "oracle/asscpp_oracle.cpp (CPU restatement, one DDA per occupied voxel)")
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * tot / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(a),
        "arm": {"cull": "brute (reference algorithm: every point tested per view)", "views_evaluated_per_step": k},
        "point_rays_per_sec": rays / tot,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": nthreads, "kind": kind,
                         "sample": "first %d viewpoints of each step's batch x %d steps of the same workload (%s, %d "
                                   "threads)" % (k, a.steps, what, nthreads)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if os.environ.get("ASSCPP_BENCH_REPORT_MAPS"):
        out["mapped_repo_libraries"] = sorted({ln.split()[-1] for ln in open("/proc/self/maps")
                                               if ".so" in ln and ROOT in ln})
    print(json.dumps(out))
    return 0


def host_memory_bytes():
    try:
        return os.sysconf("SC_PAGE_SIZE") * os.sysconf("SC_PHYS_PAGES")
    except (ValueError, OSError):
        return 64 << 30


def main():
    a = parse()
    if a.impl == "reference":
        return run_reference(a)
    import torch
    import torch.distributed as dist
    import asscpp_b200 as A

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_gpus = world

    xyz, pool = make_inputs(a, rank)
    params = A.default_params(leaf=a.leaf, cov_leaf=a.cov_leaf)
    t0 = time.perf_counter()
    ctx = A.CoverageContext(xyz, params, device=local)
    t_create = time.perf_counter() - t0
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.from_numpy(A.CoverageContext.comm_unique_id()))
        dist.broadcast(uid, 0)
        ctx.comm_init(world, rank, uid.cpu().numpy())

    B = a.views_per_step
    flags = A.EVAL_BRUTE if a.brute else 0
    total_steps = a.warmup + a.steps
    # the rank's batch twice back to back: step s reads B consecutive poses starting at its rotation offset
    pool2 = np.ascontiguousarray(np.concatenate([pool, pool]))
    d_pool = torch.from_numpy(pool2).cuda()
    h_pool = torch.from_numpy(pool2).pin_memory()
    h_np = h_pool.numpy()
    flush_buf = None if a.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def flush():
        if flush_buf is not None:
            flush_buf.zero_()
            torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.sync()

    def step(s, covered, device_poses):
        """One pass of the search loop's scoring step (see the module docstring)."""
        lo = step_offset(s, B)
        if device_poses:
            ctx.eval_ex(d_pool.data_ptr() + lo * 56, covered=covered, flags=flags, nviews=B)
        else:
            ctx.eval_ex(h_np[lo:lo + B], covered=covered, flags=flags)
        gp, gv = ctx.allgather_gain(B) if world > 1 else ctx.gains(B)
        w = best_view(gp, gv)
        if w // B == rank:
            covered.or_views(h_np[lo + w % B:lo + w % B + 1], -1)
        if world > 1:
            ctx.allreduce_or(covered)
        return w

    # ---- leg 1: device-resident poses ("value") ---------------------------------------------------
    cov1 = ctx.new_set()
    for s in range(a.warmup):
        step(s, cov1, True)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches = 0
    ms_total, kern_ms, kern_launches = 0.0, 0.0, 0
    coll = [[0.0, 0], [0.0, 0]]
    res_all = []
    barrier()
    for s in range(a.warmup, total_steps):
        flush()
        l0 = ctx.launch_count()
        ctx.timer_begin()
        step(s, cov1, True)
        ms_total += ctx.timer_end()
        launches += ctx.launch_count() - l0
        km, kn = ctx.eval_kernel_ms()
        (ag, nag), (ar, nar) = ctx.collective_ms()
        coll[0][0] += ag; coll[0][1] += nag; coll[1][0] += ar; coll[1][1] += nar
        # the accepted view's one-view launch is timed too: keep the dominant kernel's figures to the batch launch
        kern_ms += km
        kern_launches += kn
        if len(res_all) < 4:
            res_all.append(ctx_results_of_batch(ctx, A, d_pool, step_offset(s, B), B, cov1, flags))
    barrier()
    clocks = sampler.stop()
    res = np.concatenate(res_all)
    cov_points, cov_voxels = cov1.popcount()

    # ---- leg 2: host buffers through the C ABI ("e2e") ----------------------------------------------
    cov2 = ctx.new_set()
    for s in range(a.warmup):
        step(s, cov2, False)
    barrier()
    e2e_s = 0.0
    for s in range(a.warmup, total_steps):
        flush()
        barrier()
        t = time.perf_counter()
        step(s, cov2, False)
        ctx.sync()
        e2e_s += time.perf_counter() - t
    barrier()

    views_sampled = len(res)
    t_max = torch.tensor([ms_total, e2e_s * 1e3, kern_ms], dtype=torch.float64, device="cuda")
    per_view = torch.tensor([float(res["survivors"].sum(dtype=np.float64)), float(res["unique_rays"].sum(dtype=np.float64)),
                             float(res["steps"].sum(dtype=np.float64)), float(res["visible"].sum(dtype=np.float64))],
                            dtype=torch.float64, device="cuda") / views_sampled
    if world > 1:
        dist.all_reduce(t_max, op=dist.ReduceOp.MAX)
        dist.all_reduce(per_view, op=dist.ReduceOp.SUM)
        per_view /= world
    ms_total_max, e2e_ms_max, kern_ms_max = [float(x) for x in t_max.cpu()]
    surv_v, uniq_v, steps_v, vis_v = [float(x) for x in per_view.cpu()]

    if rank == 0:
        views_total = B * a.steps * world
        secs = ms_total_max * 1e-3
        value = views_total / secs
        peak, peak_src = measured_peak()
        # roofline of the dominant kernel on THIS rank: the batch launches (one per step) + the one-view launches of
        # the accepted viewpoints are all eval_views_kernel; algorithmic bytes from the sampled batches' statistics
        alg_per_view = algorithmic_bytes(a.points, res) / views_sampled
        launches_batch = a.steps
        avg_launch_s = kern_ms * 1e-3 / max(1, launches_batch)   # the one-view launches add < 0.1 %
        alg_per_launch = alg_per_view * B
        achieved = alg_per_launch / avg_launch_s / 1e9
        props = torch.cuda.get_device_properties(local)
        mhz = float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0)
        fp32_peak = props.multi_processor_count * 128 * 2 * mhz * 1e6 / 1e12
        fp32_alg = 48.0 * a.points * B / avg_launch_s / 1e12
        roof = {"bound": "hbm", "kernel": "eval_views_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": None, "peak_source": peak_src, "launches": int(kern_launches),
                "avg_launch_ms": avg_launch_s * 1e3, "algorithmic_bytes_per_launch": alg_per_launch,
                "fp32": {"achieved": fp32_alg, "peak": fp32_peak, "unit": "TFLOP/s", "frac": fp32_alg / fp32_peak,
                         "note": "reference algorithm's 48 flop per point-view plane test x N x views / launch time; "
                                 "> 1 for the same reason as frac: the bucket pre-cull skips the tests"},
                "note": "frac = contract figure (algorithmic bytes of the reference's per-view scan, SURVEY 8d); the "
                        "bucket pre-cull reads only frustum-adjacent buckets, so frac > 1 means bytes avoided, not "
                        "bandwidth.  dram = bytes the kernel really moved; issue = the bound it runs against"}
        cap = ncu_capture(a)
        if cap:
            roof["traffic"] = cap["dram_bytes_per_launch"]
            roof["dram"] = {"achieved": cap["dram_bytes_per_launch"] / avg_launch_s / 1e9, "peak": peak, "unit": "GB/s",
                            "frac": cap["dram_bytes_per_launch"] / avg_launch_s / 1e9 / peak,
                            "read_bytes": cap.get("dram_read_bytes"), "write_bytes": cap.get("dram_write_bytes"),
                            "source": cap.get("source")}
            if cap.get("warp_instructions_per_launch"):
                ipc = cap["warp_instructions_per_launch"] / avg_launch_s / (props.multi_processor_count * mhz * 1e6)
                roof["issue"] = {"bound": "issue", "achieved": ipc, "peak": 4.0, "unit": "warp inst/cycle/SM",
                                 "frac": ipc / 4.0, "warp_instructions_per_launch": cap["warp_instructions_per_launch"],
                                 "source": "instruction count from %s (a capture of this workload, not this run); "
                                           "time and clock from this run" % cap.get("source")}
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(a),
            "arm": {"cull": "brute" if a.brute else "bucketed",
                    "parallelism": "viewpoints sharded over %d GPU(s), points and covered set replicated" % n_gpus,
                    "l2": "L2 flushed between timed steps (256 MiB write); point SoA (16 B x N) exceeds L2"},
            "point_rays_per_sec": surv_v * value, "unique_rays_per_sec": uniq_v * value,
            "dda_steps_per_sec": steps_v * value,
            "mean_survivors_per_view": surv_v, "mean_unique_rays_per_view": uniq_v, "mean_visible_per_view": vis_v,
            "covered_after_run": {"points": cov_points, "coverage_voxels": cov_voxels,
                                  "model_coverage_voxels": ctx.n_cov_voxels},
            "e2e": {"value": views_total / (e2e_ms_max * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": 56 * B + 56, "d2h_bytes_per_step": 8 * B * world + 16},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roof,
            "collective_ms": None if world == 1 else {
                "allgather_gain": {"ms_per_step": coll[0][0] / a.steps, "bytes_per_rank": 8 * B, "calls": coll[0][1]},
                "allreduce_or": {"ms_per_step": coll[1][0] / a.steps,
                                 "bytes_per_rank": (a.points + 7) // 8 + (ctx.n_cov_voxels + 7) // 8, "calls": coll[1][1]},
                "note": "device time on rank 0 between CUDA events around each collective (includes waiting for the slowest rank)"},
            "create_seconds": t_create,
        }
        if n_gpus == 1 and not a.no_cpu_baseline:
            from oracle import oracle as O
            nthreads = os.cpu_count() or 1
            orc = O.Oracle(xyz, O.make_params(leaf=a.leaf, cov_leaf=a.cov_leaf))
            cp, cv = cov1.export()
            sample_poses = pool2[step_offset(a.warmup, B):step_offset(a.warmup, B) + B]
            cpu = cpu_sample(orc, O, sample_poses, a.cpu_seconds, nthreads, (cp, cv))
            cpu1 = cpu_sample(orc, O, sample_poses, a.cpu_seconds * 0.6, 1, (cp, cv))
            out["cpu_baseline"] = {"value": cpu["ve_per_s"], "unit": UNIT, "cores": nthreads, "kind": "port",
                                   "point_rays_per_sec": cpu["point_rays_per_s"],
                                   "single_thread": {"value": cpu1["ve_per_s"], "point_rays_per_sec": cpu1["point_rays_per_s"],
                                                     "views": cpu1["views"], "seconds": cpu1["seconds"]},
                                   "sample": "first %d viewpoints of the first timed step, oracle/asscpp_oracle.cpp "
                                             "(one DDA per occupied voxel) on %d host threads, %.1f s; single thread: "
                                             "first %d viewpoints, %.1f s"
                                             % (cpu["views"], nthreads, cpu["seconds"], cpu1["views"], cpu1["seconds"])}
        print(json.dumps(out))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def ctx_results_of_batch(ctx, A, d_pool, lo, B, covered, flags):
    """Full per-view records of one batch (survivors, cells, DDA steps: the roofline's statistics), fetched outside
    the timed region: the timed step only moves the 8-byte scores."""
    ctx.eval_ex(d_pool.data_ptr() + lo * 56, covered=covered, flags=flags, nviews=B)
    return ctx.results(B)


if __name__ == "__main__":
    sys.exit(main())
