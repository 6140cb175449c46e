"""Research instrument (not product, not oracle): how many DDA steps of a view run through empty B^3 bricks?
Approximate float64 DDA, statistics only."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import asscpp_b200 as A
from oracle import oracle as O

N = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
NV = int(sys.argv[2]) if len(sys.argv) > 2 else 12
leaf = 0.05
xyz = A.synth_structure(N)
poses = A.synth_views(400)
o = O.Oracle(xyz, O.make_params(leaf=leaf, cov_leaf=0.5))
rng = np.random.default_rng(0)
tot = {}
done = 0
for v in rng.permutation(len(poses)):
    r = o.eval_view(poses[v])
    surv = r["survivors"]
    if len(surv) < 1000:
        continue
    P = xyz[surv]
    ijk = np.floor(P * np.float32(1 / leaf)).astype(np.int64)
    mn, mx = ijk.min(0), ijk.max(0)
    div = mx - mn + 1
    occ = np.zeros(div + 2, bool)  # halo of one
    rel = ijk - mn + 1
    occ[rel[:, 0], rel[:, 1], rel[:, 2]] = True
    vox = np.unique(rel, axis=0)
    origin = poses[v][:3] / leaf - mn + 1  # in voxel units, halo frame
    ctr = vox + 0.5
    d = ctr - origin
    d /= np.linalg.norm(d, axis=1)[:, None]
    # entry into the box [1, div+1]
    with np.errstate(divide="ignore", invalid="ignore"):
        t0 = (np.where(d >= 0, 1.0, div + 1.0) - origin) / d
        t1 = (np.where(d >= 0, div + 1.0, 1.0) - origin) / d
    tmin = np.nanmax(t0, axis=1)
    start = origin + tmin[:, None] * d
    cur = np.floor(start + 0.5).astype(np.int64)  # round-to-nearest start voxel (quirk Q7)
    step = np.where(d >= 0, 1, -1)
    nxt = cur + 0.5 + 0.5 * step
    with np.errstate(divide="ignore", invalid="ignore"):
        tmax = tmin[:, None] + (nxt - start) / d
        tdel = 1.0 / np.abs(d)
    alive = np.ones(len(vox), bool)
    stats = {}
    for B in (4, 8):
        nb = (div + 2 + B - 1) // B
        pad = np.zeros(nb * B, bool); pad[:div[0] + 2, :div[1] + 2, :div[2] + 2] = occ
        bo = pad.reshape(nb[0], B, nb[1], B, nb[2], B).any(axis=(1, 3, 5))
        stats[B] = {"bo": bo, "empty_steps": 0, "runs": 0, "prev": np.full((len(vox), 3), -1)}
    steps = 0
    it = 0
    while alive.any() and it < 2000:
        it += 1
        idx = np.nonzero(alive)[0]
        c = cur[idx]
        inside = ((c >= 0) & (c <= div + 1)).all(1)
        alive[idx[~inside]] = False
        idx, c = idx[inside], c[inside]
        steps += len(idx)
        at_target = (c == vox[idx]).all(1)
        o_here = occ[c[:, 0], c[:, 1], c[:, 2]]
        dd = ((c - vox[idx]) ** 2).sum(1)
        blocked = o_here & (dd > 4) & ~at_target
        for B in (4, 8):
            s = stats[B]
            b = c // B
            e = ~s["bo"][b[:, 0], b[:, 1], b[:, 2]]
            s["empty_steps"] += int(e.sum())
            newrun = e & (b != s["prev"][idx]).any(1)
            s["runs"] += int(newrun.sum())
            s["prev"][idx] = np.where(e[:, None], b, -1)
        stop = at_target | blocked
        alive[idx[stop]] = False
        idx = idx[~stop]
        tm = tmax[idx]
        ax = np.where((tm[:, 0] <= tm[:, 1]) & (tm[:, 0] <= tm[:, 2]), 0, np.where(tm[:, 1] <= tm[:, 2], 1, 2))
        r_ = np.arange(len(idx))
        tmax[idx, ax] += tdel[idx, ax]
        cur[idx, ax] += step[idx, ax]
    print("view %d S %d U %d steps %d (%.1f/ray) | B4 empty %.1f%% runs %d (%.1f steps/run) | B8 empty %.1f%% runs %d (%.1f steps/run)" % (
        v, len(surv), len(vox), steps, steps / len(vox),
        100 * stats[4]["empty_steps"] / steps, stats[4]["runs"], stats[4]["empty_steps"] / max(1, stats[4]["runs"]),
        100 * stats[8]["empty_steps"] / steps, stats[8]["runs"], stats[8]["empty_steps"] / max(1, stats[8]["runs"])), flush=True)
    done += 1
    if done >= NV:
        break
