#!/usr/bin/env python
"""bench.py -- viewpoint-coverage evaluation throughput (BASELINE.json metric) on N B200s.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --steps K --warmup W    # the reference algorithm on the host cores

Workload (config.workload): BASELINE.json configs[2] sized for one GPU -- synthetic 10M-point
structure (synth_structure_v1), 100k candidate viewpoints (synth_views_v1), 0.05 m occlusion grid.
A "step" is one batch of --views-per-step viewpoints per GPU pushed through the hot path
(frustum cull -> per-view bit-grid -> 3D-DDA -> gain).  N > 1: viewpoints are sharded over the
ranks (weak scaling: every rank gets its own batch), the point SoA is replicated, and the only
collective is the gain-score allgather (NCCL) at the end of each step.

`value`  : VE/s with poses resident in HBM, device-timed (CUDA events on the launching stream),
           max over ranks.
`e2e`    : same metric through the host-buffer C-ABI call (asscpp_eval_batch + asscpp_get_results),
           H2D of the poses and D2H of the per-view scores inside the timed region.
`roofline`: for the dominant kernel (eval_views_kernel); algorithmic bytes per VE follow SURVEY.md 8(d):
           B_VE = 16 N + 20 S_v + G_v/8 + 2 N/8 + 36.  The kernel pre-culls voxel buckets, so it moves far
           fewer bytes than the reference algorithm nominally touches: frac > 1 is expected and
           `traffic` (ncu dram bytes per launch) is the figure to read next to it.
           `roofline.issue` (default workload only) is the bound the kernel actually runs against: warp
           instructions per launch from the committed ncu capture / this run's launch time / (SMs x sampled SM
           clock), against 4 per cycle per SM.
"""
import argparse
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--points", type=int, default=10_000_000)
    ap.add_argument("--views", type=int, default=100_000, help="size of the candidate viewpoint pool")
    ap.add_argument("--views-per-step", type=int, default=10_000)
    ap.add_argument("--leaf", type=float, default=0.05)
    ap.add_argument("--cov-leaf", type=float, default=0.5)
    ap.add_argument("--brute", action="store_true", help="disable the bucket pre-cull (stream all points per view)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-flush", action="store_true")
    return ap.parse_args()


def workload_config(a, n_gpus):
    return {
        "workload": "synth_structure_v1 %d points x %d-view pool, leaf %.3g m, cov leaf %.3g m, frustum 58x45 deg 0.7-6 m"
                    % (a.points, a.views, a.leaf, a.cov_leaf),
        "baseline_config": "configs[2] sized for one GPU (10M points / 100k viewpoints / 0.05 m)",
        "points": a.points, "view_pool": a.views, "views_per_step_per_gpu": a.views_per_step,
        "leaf_m": a.leaf, "cov_leaf_m": a.cov_leaf, "cull": "brute" if a.brute else "bucketed",
        "parallelism": "viewpoints sharded over %d GPU(s), points replicated" % n_gpus,
        "l2": "L2 flushed between timed steps (256 MiB write); point SoA (16 B x N) exceeds L2",
    }


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


def ncu_traffic():
    """dram bytes per eval_views_kernel launch from the committed ncu capture, if any."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("eval_views_kernel_dram_bytes_per_launch")
        except Exception:
            return None
    return None


def issue_roofline(kern_ms, kern_launches, clocks, views_per_launch, n_points, device):
    """What actually bounds eval_views_kernel: warp-instruction issue.  Instructions per launch come from the
    committed ncu capture (valid for the default workload only); rate = that count / the launch time measured in
    this run / (SMs x SM clock sampled in this run); peak = 4 warp instructions per cycle per SM."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        import torch
        t = json.load(open(p))
        inst = t.get("eval_views_kernel_warp_instructions_per_launch")
        if not inst or t.get("views_per_launch") != views_per_launch or t.get("points") != n_points:
            return None
        sms = torch.cuda.get_device_properties(device).multi_processor_count
        mhz = float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 0.0)
        if mhz <= 0 or kern_launches <= 0:
            return None
        ipc = inst / (kern_ms / kern_launches * 1e-3) / (sms * mhz * 1e6)
        return {"bound": "issue", "achieved": ipc, "peak": 4.0, "unit": "warp inst/cycle/SM", "frac": ipc / 4.0,
                "warp_instructions_per_launch": inst, "source": "profiles/r01_ncu_eval_views_full.txt"}
    except Exception:
        return None


def algorithmic_bytes(n_points, res):
    """B_VE = 16 N + 20 S_v + G_v/8 + 2 N/8 + 36 summed over the views."""
    v = len(res)
    return (16.0 * n_points + 2.0 * n_points / 8.0 + 36.0) * v + 20.0 * float(res["survivors"].sum(dtype=np.float64)) \
        + float(res["cells"].sum(dtype=np.float64)) / 8.0


def cpu_leg(a, xyz, poses, seconds, nthreads):
    """The oracle (CPU restatement of the reference algorithm) on the host cores: bounded sample."""
    from oracle import oracle as O
    orc = O.Oracle(xyz, O.make_params(leaf=a.leaf, cov_leaf=a.cov_leaf))
    probe = poses[:nthreads]
    t = time.perf_counter()
    orc.eval_batch(probe, memo=True, nthreads=nthreads)
    per_round = max(time.perf_counter() - t, 1e-4)
    k = int(max(nthreads, min(len(poses), nthreads * max(1, round(seconds / per_round)))))
    sample = poses[:k]
    t = time.perf_counter()
    st = orc.eval_batch(sample, memo=True, nthreads=nthreads)
    dt = time.perf_counter() - t
    return orc, {"views": k, "seconds": dt, "ve_per_s": k / dt,
                 "point_rays_per_s": float(st[:, O.ST_SURVIVORS].sum()) / dt}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import asscpp_b200 as A
    nthreads = os.cpu_count() or 1
    xyz = A.synth_structure(a.points)
    poses = A.synth_views(a.views)
    from oracle import oracle as O
    orc = O.Oracle(xyz, O.make_params(leaf=a.leaf, cov_leaf=a.cov_leaf))
    # bounded sample per step: about cpu_seconds of work over all steps
    t = time.perf_counter()
    orc.eval_batch(poses[:nthreads], memo=True, nthreads=nthreads)
    per_round = max(time.perf_counter() - t, 1e-4)
    total_steps = a.steps + a.warmup
    rounds = max(1, int(round(max(a.cpu_seconds, 10.0) * 4 / total_steps / per_round)))
    k = min(nthreads * rounds, a.views_per_step)
    times, rays = [], 0.0
    for s in range(total_steps):
        lo = (s * k) % max(1, len(poses) - k)
        t = time.perf_counter()
        st = orc.eval_batch(poses[lo:lo + k], memo=True, nthreads=nthreads)
        dt = time.perf_counter() - t
        if s >= a.warmup:
            times.append(dt)
            rays += float(st[:, O.ST_SURVIVORS].sum())
    tot = sum(times)
    value = k * a.steps / tot
    cfg = workload_config(a, 1)
    cfg["cull"] = "brute (reference algorithm: every point tested per view)"
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * tot / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
        "point_rays_per_sec": rays / tot,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": nthreads, "kind": "port",
                         "sample": "%d views per step x %d steps of the same workload (oracle/asscpp_oracle.cpp, "
                                   "memoised per voxel, %d threads)" % (k, a.steps, nthreads)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "the reference itself (PCL/ROS) cannot be built in this image; this is the CPU port of its algorithm",
    }
    print(json.dumps(out))
    return 0


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

    xyz = A.synth_structure(a.points)
    pool = A.synth_views(a.views)
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

    def step_views(s):
        # global batch of world*B views per step, rank r takes the r-th slice; the pool is cycled
        lo = ((s * world + rank) * B) % max(1, len(pool) - B + 1)
        return pool[lo:lo + B]

    d_pose_steps = [torch.from_numpy(np.ascontiguousarray(step_views(s))).cuda() for s in range(total_steps)]
    h_pose_steps = [torch.from_numpy(np.ascontiguousarray(step_views(s))).pin_memory() for s in range(total_steps)]
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

    def device_step(s):
        ctx.eval_device(d_pose_steps[s].data_ptr(), B, None, flags)
        if world > 1:
            ctx.allgather_results_device(B)

    def host_step(s):
        ctx.eval_async(h_pose_steps[s].numpy(), None, flags)
        if world > 1:
            return ctx.allgather_results(B)
        return ctx.results(B)

    # ---- leg 1: device-resident poses ("value") ---------------------------------------------------
    for s in range(a.warmup):
        device_step(s)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = ctx.launch_count()
    ms_total, kern_ms, kern_launches = 0.0, 0.0, 0
    res_all = []
    barrier()
    for s in range(a.warmup, total_steps):
        flush()
        ctx.timer_begin()
        device_step(s)
        ms_total += ctx.timer_end()
        km, kn = ctx.eval_kernel_ms()
        kern_ms += km
        kern_launches += kn
        res_all.append(ctx.results(B))  # outside the timed region (stats for the roofline)
    barrier()
    launches = ctx.launch_count() - launches0
    clocks = sampler.stop()
    res = np.concatenate(res_all)

    # ---- leg 2: host buffers through the C ABI ("e2e") ----------------------------------------------
    for s in range(a.warmup):
        host_step(s)
    barrier()
    e2e_s = 0.0
    for s in range(a.warmup, total_steps):
        flush()
        barrier()
        t = time.perf_counter()
        host_step(s)
        e2e_s += time.perf_counter() - t
    barrier()

    t_max = torch.tensor([ms_total, e2e_s * 1e3, kern_ms], dtype=torch.float64, device="cuda")
    sums = torch.tensor([float(res["survivors"].sum(dtype=np.float64)), float(res["unique_rays"].sum(dtype=np.float64)),
                         float(res["steps"].sum(dtype=np.float64)), float(res["visible"].sum(dtype=np.float64)),
                         algorithmic_bytes(a.points, res)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_max, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    ms_total_max, e2e_ms_max, kern_ms_max = [float(x) for x in t_max.cpu()]
    surv, uniq, dda_steps, vis, alg_bytes_all = [float(x) for x in sums.cpu()]

    if rank == 0:
        views_total = B * a.steps * world
        value = views_total / (ms_total_max * 1e-3)
        peak, peak_src = measured_peak()
        # roofline of the dominant kernel on THIS rank's launches
        alg_bytes = algorithmic_bytes(a.points, res)
        achieved = alg_bytes / (kern_ms * 1e-3) / 1e9
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(a, n_gpus),
            "point_rays_per_sec": surv / (ms_total_max * 1e-3),
            "unique_rays_per_sec": uniq / (ms_total_max * 1e-3),
            "dda_steps_per_sec": dda_steps / (ms_total_max * 1e-3),
            "mean_survivors_per_view": surv / views_total, "mean_unique_rays_per_view": uniq / views_total,
            "mean_visible_per_view": vis / views_total,
            "e2e": {"value": views_total / (e2e_ms_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 56 * B,
                    "d2h_bytes_per_step": 64 * B * (world if world > 1 else 1)},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "eval_views_kernel", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": ncu_traffic(),
                         "peak_source": peak_src, "launches": int(kern_launches),
                         "avg_launch_ms": kern_ms / max(1, kern_launches),
                         "algorithmic_bytes_per_launch": alg_bytes / max(1, kern_launches),
                         "note": "algorithmic bytes = reference algorithm's per-view scan (SURVEY 8d); the bucket "
                                 "pre-cull reads only frustum-adjacent buckets, so frac > 1 means bytes avoided, "
                                 "not bandwidth; see traffic and DESIGN.md"},
            "create_seconds": t_create,
        }
        iss = issue_roofline(kern_ms, kern_launches, clocks, B, a.points, local)
        if iss is not None and not a.brute:
            out["roofline"]["issue"] = iss
        if n_gpus == 1 and not a.no_cpu_baseline:
            nthreads = os.cpu_count() or 1
            _, cpu = cpu_leg(a, xyz, step_views(a.warmup), a.cpu_seconds, nthreads)
            out["cpu_baseline"] = {"value": cpu["ve_per_s"], "unit": UNIT, "cores": nthreads, "kind": "port",
                                   "point_rays_per_sec": cpu["point_rays_per_s"],
                                   "sample": "first %d views of the first timed step, oracle/asscpp_oracle.cpp "
                                             "(memoised per voxel) on %d host threads, %.1f s"
                                             % (cpu["views"], nthreads, cpu["seconds"])}
        print(json.dumps(out))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
