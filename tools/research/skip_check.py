"""Empty-space skip of the DDA: every skip replayed step by step on the device (EVAL_VERIFY_SKIP) and the whole
result compared with the plain walk (EVAL_NO_SKIP).  Usage: python tools/skip_check.py [n_points] [n_views]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import asscpp_b200 as A  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
nv = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000
bad = 0
for leaf, cov in ((0.05, 0.5), (0.1, 0.5), (0.5, 0.5)):
    xyz = A.synth_structure(n)
    poses = A.synth_views(nv)
    ctx = A.CoverageContext(xyz, A.default_params(leaf=leaf, cov_leaf=cov))
    ctx.phase_cycles(True)
    a = ctx.eval_batch(poses, flags=A.EVAL_HASH | A.EVAL_VERIFY_SKIP)
    pc = ctx.phase_cycles(True)
    b = ctx.eval_batch(poses, flags=A.EVAL_HASH | A.EVAL_NO_SKIP)
    ctx.phase_cycles(False)
    diff = sum(int((a[f] != b[f]).sum()) for f in a.dtype.names if f not in ("status", "reserved"))
    print("leaf %.2f: attempts %d ok %d (%.1f%%) skipped steps %d of %d (%.1f%%), %.1f steps/skip, replay mismatches %d, "
          "result fields differing from the plain walk %d" % (
              leaf, pc[13], pc[14], 100.0 * pc[14] / max(1, pc[13]), pc[15], int(a["steps"].sum()),
              100.0 * pc[15] / max(1, int(a["steps"].sum())), pc[15] / max(1, pc[14]), pc[12], diff), flush=True)
    bad += int(pc[12]) + diff
    for flags, name in ((0, "skip"), (A.EVAL_NO_SKIP, "plain")):
        ctx.eval_batch(poses[:2000], flags=flags)
        t = time.perf_counter()
        ctx.eval_batch(poses, flags=flags)
        dt = time.perf_counter() - t
        print("   %-5s %.2f ms  %.0f VE/s" % (name, dt * 1e3, len(poses) / dt), flush=True)
    ctx.close()
print("TOTAL BAD", bad)
sys.exit(1 if bad else 0)
