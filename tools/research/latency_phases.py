"""Where the time of a SMALL batch goes (the A* use): per-phase SM cycles per view on the bundled model, and the size
of the work (survivors, rays, visible points) of those views."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import asscpp_b200 as A
g = np.load(os.path.join(ROOT, "tests", "golden", "etihad_model.npz"))
sp = np.load(os.path.join(ROOT, "tests", "golden", "etihad_searchspace_2p5m.npz"))
ctx = A.CoverageContext(g["xyz"], A.default_params())
s = ctx.new_set(); s.or_views(sp["cam0"][600:620], 5)

print("model: points %d occ voxels %d cov voxels %d" % (ctx.n_points, ctx.n_occ_voxels if hasattr(ctx, "n_occ_voxels") else -1, ctx.n_cov_voxels))
for n in (2, 40, 400):
    poses = np.stack([sp["cam0"][:n // 2], sp["cam1"][:n // 2]], 1).reshape(-1, 7)
    for _ in range(5): ctx.eval_ex(poses, covered=s, group_size=2); ctx.gains(n)
    ctx.phase_cycles(True)
    reps = 20
    for _ in range(reps): ctx.eval_ex(poses, covered=s, group_size=2); ctx.gains(n)
    ph = ctx.phase_cycles(False)
    res = ctx.results(n)
    views = max(int(ph[4]), 1)
    tot = float(ph[0] + ph[1] + ph[2] + ph[3])
    print("views %4d: phase cycles per non-empty view: cull %.0f grid %.0f dda %.0f visit %.0f (sum %.0f = %.1f us at 1.9 GHz); "
          "non-empty views %d of %d per rep; CTA lifetime cycles (sum over CTAs, per rep) %.0f" % (
              n, ph[0] / views, ph[1] / views, ph[2] / views, ph[3] / views, tot / views, tot / views / 1900.0,
              views // reps, n, ph[5] / reps))
    print("   cull split per view: P1a %.0f P1b %.0f P2 classify %.0f P2 cut points %.0f; cut voxels(warp0) %.1f candidates %.0f" % (
        ph[6] / views, ph[7] / views, ph[8] / views, ph[9] / views, ph[10] / views, ph[11] / views))
    print("   reset+planes %.0f; visit pass A (lane per voxel) %.0f of the visit; warp 0 waits %.0f for the other warps at the end of P2; group gains %.0f" % (
        ph[12] / views, ph[13] / views, ph[14] / views, ph[15] / views))
    ne = res["survivors"] > 0
    print("   work per non-empty view: survivors %.0f rays %.0f visible %.0f steps %.0f cells %.0f; max survivors %d max rays %d" % (
        res["survivors"][ne].mean() if ne.any() else 0, res["unique_rays"][ne].mean() if ne.any() else 0,
        res["visible"][ne].mean() if ne.any() else 0, res["steps"][ne].mean() if ne.any() else 0,
        res["cells"][ne].mean() if ne.any() else 0, res["survivors"].max(), res["unique_rays"].max()))
