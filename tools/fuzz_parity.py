"""Randomised parity run: GPU path against the oracle over random model sizes, lattices, frusta and PCD origins.
Every integer field and both set hashes per view, visible lists of a few views, grouped gains, accuracy extrema,
uncovered clusters.  Usage: python tools/fuzz_parity.py [seconds] [seed]"""
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

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
threads = os.cpu_count() or 1
t_end = time.time() + budget
bad = cases = views_total = 0
while time.time() < t_end:
    n = int(rng.choice([20_000, 100_000, 400_000, 1_500_000]))
    leaf = float(rng.choice([0.03, 0.05, 0.08, 0.1, 0.25, 0.5, 0.6]))
    cov = float(rng.choice([leaf, 0.5, 0.3]))
    hfov, vfov = float(rng.uniform(30, 100)), float(rng.uniform(25, 80))
    near = float(rng.uniform(0.1, 1.5))
    far = float(near + rng.uniform(1.0, 6.0))
    if leaf < 0.05:
        far = min(far, near + 3.0)  # keeps the per-view grid inside the scratch bound of both sides
    origin = tuple(float(x) for x in (rng.uniform(-5, 45, 3) if rng.random() < 0.5 else np.zeros(3)))
    seed = int(rng.integers(1, 2 ** 31))
    xyz = A.synth_structure(n, point_seed=seed)
    poses = A.synth_views(160, view_seed=seed ^ 0x5555)
    kw = dict(leaf=leaf, cov_leaf=cov, hfov_deg=hfov, vfov_deg=vfov, near_dist=near, far_dist=far, pcd_origin=origin)
    ctx = A.CoverageContext(xyz, A.default_params(**kw))
    orc = O.Oracle(xyz, O.make_params(leaf=leaf, cov_leaf=cov, hfov=hfov, vfov=vfov, near=near, far=far, pcd_origin=origin))
    res = ctx.eval_batch(poses, flags=A.EVAL_HASH)
    st = orc.eval_batch(poses, memo=True, nthreads=threads)
    case_bad = 0
    for name, col in FIELDS:
        case_bad += int((res[name].astype(np.uint64) != st[:, col].astype(np.uint64)).sum())
    for v in np.nonzero(st[:, O.ST_VISIBLE] > 0)[0][:2]:
        case_bad += int(not np.array_equal(ctx.get_visible(poses[v]), orc.eval_view(poses[v])["visible"]))
    # covered set + grouped gains
    s = ctx.new_set()
    s.or_views(poses[:40], 5)
    cv = np.zeros(orc.vox_words(), np.uint64)
    cp = np.zeros(orc.pts_words(), np.uint64)
    orc.accumulate(poses[:40], cp, cv, min_visible=5)
    gp, gv = s.export()
    case_bad += int(not np.array_equal(gp, cp)) + int(not np.array_equal(gv, cv))
    in_cov = np.unpackbits(cv.view(np.uint8), bitorder="little").astype(bool)
    grp = ctx.eval_groups(poses[40:100], 3, covered=s)["gain_vox"].astype(np.int64).reshape(-1, 3).sum(1)
    ids = orc.cov_ids()
    for gi in range(20):
        u = np.unique(np.concatenate([ids[orc.eval_view(p)["visible"]] for p in poses[40 + 3 * gi: 43 + 3 * gi]]).astype(np.int64))
        case_bad += int(grp[gi] != int((~in_cov[u]).sum()))
    # uncovered clusters
    ids_u, ctr_u = orc.uncovered(cv)
    tol = 1.6 * cov  # heuristic.cpp:160 uses 0.8 at the 0.5 m lattice; the device supports up to 3 voxels of reach
    lab_u, k_u, _ = O.euclidean_clusters(ctr_u, tol, 5, 50000)
    g_ids, _ = s.uncovered()
    _, cl = s.uncovered_clusters(tol, 5, 50000)
    case_bad += int(not np.array_equal(g_ids, ids_u)) + int(len(cl) != k_u)
    cases += 1
    views_total += len(poses)
    bad += case_bad
    print("case %d: n %d leaf %.2f cov %.2f fov %.0fx%.0f range %.2f-%.2f origin %s: visible %d, carry views %d, mismatches %d"
          % (cases, n, leaf, cov, hfov, vfov, near, far, "zero" if origin == (0.0, 0.0, 0.0) else "random",
             int(res["visible"].sum()), int((res["carry"] > 0).sum()), case_bad), flush=True)
    ctx.close()
print("FUZZ cases %d views %d TOTAL MISMATCHES %d" % (cases, views_total, bad))
sys.exit(1 if bad else 0)
