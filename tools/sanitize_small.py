"""Small run over every kernel family for compute-sanitizer (memcheck / racecheck):
   compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import asscpp_b200 as A  # noqa: E402

g = np.load(os.path.join(ROOT, "tests", "golden", "etihad_model.npz"))
sp = np.load(os.path.join(ROOT, "tests", "golden", "etihad_searchspace_2p5m.npz"))
ms = np.load(os.path.join(ROOT, "tests", "golden", "etihad_mesh.npz"))
ctx = A.CoverageContext(g["xyz"], A.default_params())
poses = np.stack([sp["cam0"][600:640], sp["cam1"][600:640]], 1).reshape(-1, 7)
r = ctx.eval_batch(poses, flags=A.EVAL_HASH | A.EVAL_ACCURACY)
s = ctx.new_set()
s.or_views(poses[:20], 5)
k = ctx.eval_groups(poses, 2, covered=s)
acc = ctx.new_set()
n = ctx.eval_ex(poses, covered=s, accum=acc, accum_min_visible=5)
gp, gv = ctx.gains(n)
ids, ctr = s.uncovered()
lab, cl = s.uncovered_clusters(0.8, 5, 50000)
b = ctx.eval_batch(poses[:16], flags=A.EVAL_BRUTE | A.EVAL_FORCE_GLOBAL)
xyz = A.synth_structure(200_000)
c2 = A.CoverageContext(xyz, A.default_params(leaf=0.05, cov_leaf=0.5))
v = A.synth_views(64)
r2 = c2.eval_groups(v, 4)
mesh = A.Mesh(ms["vertices"], ms["triangles"])
h, d, ok = mesh.filter(sp["uav"][:128, :3])
f = mesh.segments_free(sp["uav"][:128, :3], sp["uav"][1:129, :3])
print("ok", int(r["visible"].sum()), int(k["gain_vox"].sum()), len(ids), len(cl), int(r2["visible"].sum()), int(ok.sum()), int(f.sum()))
