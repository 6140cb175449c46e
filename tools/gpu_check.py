"""Developer check: GPU path vs CPU oracle on the bundled model and a small synthetic case.
Prints per-field mismatch counts (tests/ holds the real parity suite)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import asscpp_b200 as A  # noqa: E402
from oracle import oracle as O  # noqa: E402

FIELDS = [("survivors", O.ST_SURVIVORS), ("visible", O.ST_VISIBLE), ("carry", O.ST_CARRY),
          ("unique_rays", O.ST_UNIQUE), ("cells", O.ST_CELLS), ("steps", O.ST_STEPS),
          ("gain_pts", O.ST_GAIN_PTS), ("gain_vox", O.ST_GAIN_VOX), ("vis_hash", O.ST_HASH),
          ("surv_hash", O.ST_SURV_HASH)]


def compare(res, st, label):
    bad_total = 0
    for name, col in FIELDS:
        a = res[name].astype(np.uint64)
        b = st[:, col].astype(np.uint64)
        bad = np.nonzero(a != b)[0]
        if bad.size:
            bad_total += bad.size
            k = bad[0]
            print("  %s: %s mismatches=%d first view %d gpu=%d cpu=%d" % (label, name, bad.size, k, a[k], b[k]))
    print("%s: %d views, mismatching fields: %d, status hist %s" % (
        label, len(res), bad_total, np.unique(res["status"], return_counts=True)))
    return bad_total


def run(xyz, poses, leaf, cov_leaf, label, flags=A.EVAL_HASH):
    ctx = A.CoverageContext(xyz, A.default_params(leaf=leaf, cov_leaf=cov_leaf))
    orc = O.Oracle(xyz, O.make_params(leaf=leaf, cov_leaf=cov_leaf))
    t = time.time()
    res = ctx.eval_batch(poses, flags=flags)
    t_gpu = time.time() - t
    t = time.time()
    st = orc.eval_batch(poses, memo=True, nthreads=os.cpu_count())
    t_cpu = time.time() - t
    print("%s: gpu %.3fs (%.0f VE/s) cpu %.3fs (%.0f VE/s)" % (label, t_gpu, len(poses) / t_gpu, t_cpu, len(poses) / t_cpu))
    bad = compare(res, st, label)
    for fl, nm in ((flags | A.EVAL_BRUTE, "brute"), (flags | A.EVAL_FORCE_GLOBAL, "global-grid")):
        r2 = ctx.eval_batch(poses, flags=fl)
        bad += compare(r2, st, label + "/" + nm)
    # visible lists on a few views
    nz = np.nonzero(st[:, O.ST_VISIBLE] > 0)[0][:5]
    for v in nz:
        a = ctx.get_visible(poses[v])
        b = orc.eval_view(poses[v])["visible"]
        if not np.array_equal(a, b):
            bad += 1
            print("  visible list mismatch at view", v, len(a), len(b))
    ctx.close()
    return bad


def main():
    g = os.path.join(ROOT, "tests", "golden")
    m = np.load(os.path.join(g, "etihad_model.npz"))
    ss = np.load(os.path.join(g, "etihad_searchspace_2p5m.npz"))
    poses = np.concatenate([ss["cam0"], ss["cam1"]])
    bad = run(m["xyz"], poses, 0.5, 0.5, "etihad/0.5")
    xyz = A.synth_structure(300_000)
    pv = A.synth_views(400)
    bad += run(xyz, pv, 0.5, 0.5, "synth300k/0.5")
    bad += run(xyz, pv, 0.05, 0.5, "synth300k/0.05+cov0.5")
    bad += run(xyz, pv, 0.05, 0.05, "synth300k/0.05")
    print("TOTAL MISMATCHES", bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
