"""Developer aid: per-phase SM-cycle split of eval_views_kernel on the benchmark workload."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import asscpp_b200 as A  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
leaf = float(sys.argv[2]) if len(sys.argv) > 2 else 0.05
nv = int(sys.argv[3]) if len(sys.argv) > 3 else 10_000
xyz = A.synth_structure(n)
poses = A.synth_views(nv)
ctx = A.CoverageContext(xyz, A.default_params(leaf=leaf))
for flags, name in ((0, "bucketed"), (A.EVAL_FORCE_GLOBAL, "global-grid")):
    ctx.eval_batch(poses, flags=flags)
    ctx.phase_cycles(True)
    ctx.timer_begin()
    ctx.eval_async(poses, flags=flags)
    ms = ctx.timer_end()
    res = ctx.results(nv)
    c = ctx.phase_cycles(False).astype(np.float64)
    tot = c[:4].sum()
    print("  cull split: P1a supers %.1f%% P1b chunks %.1f%% P2 points %.1f%% of all phase cycles" % (100 * c[6] / tot, 100 * c[7] / tot, 100 * (c[0] - c[6] - c[7]) / tot))
    print("  P2 warp 0: classify loop %.1f%% cut-voxel points %.1f%% (rest = bbox/atomics/barrier wait); per view: candidates %.0f, cut voxels of warp 0 %.1f"
          % (100 * c[8] / tot, 100 * c[9] / tot, c[11] / max(c[4], 1), c[10] / max(c[4], 1)))
    print("  warp 0 at the end of P2: waits %.1f%% of all phase cycles for the other warps; reset %.1f%%; visit pass A %.1f%%"
          % (100 * c[14] / tot, 100 * c[12] / tot, 100 * c[13] / tot))
    print("%s: %.2f ms for %d views (%.0f VE/s); phases cull %.1f%% grid %.1f%% dda %.1f%% visit %.1f%%; "
          "busy/lifetime %.2f; cycles/view %.0f; smem views %d global %d; mean S %.0f U %.0f vis %.0f steps/ray %.1f words(mean) %.0f"
          % (name, ms, nv, nv / ms * 1e3, 100 * c[0] / tot, 100 * c[1] / tot, 100 * c[2] / tot, 100 * c[3] / tot,
             tot / max(c[5], 1), tot / max(c[4], 1), int((res["status"] & 1 != 0).sum()), int((res["status"] & 2 != 0).sum()),
             res["survivors"].mean(), res["unique_rays"].mean(), res["visible"].mean(),
             res["steps"].sum() / max(1, res["unique_rays"].sum()), res["cells"].mean() / 32))
ctx.close()
