"""Config 1 at the size of the reference's own search space (sspp/resources/SearchSpace*_1.5m_*: 17 583 waypoints, two
sensors each): connect the nodes (radius 2.5 m, mesh collision test on the GPU), run the SurfaceCoverageH A* with one
scoring launch per expansion, and compare with the same search scored by the CPU oracle one child at a time."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import asscpp_b200 as A  # noqa: E402
from asscpp_b200 import search as SR  # noqa: E402
from backends import OracleScorer  # noqa: E402
from oracle import oracle as O  # noqa: E402

g = os.path.join(ROOT, "tests", "golden")
xyz = np.load(os.path.join(g, "etihad_model.npz"))["xyz"]
sp = np.load(os.path.join(g, "etihad_searchspace_1p5m.npz"))
ms = np.load(os.path.join(g, "etihad_mesh.npz"))
mesh = A.Mesh(ms["vertices"], ms["triangles"])
t = time.perf_counter()
space = SR.SearchSpace.from_poses(sp["uav"], np.stack([sp["cam0"], sp["cam1"]], 1))
t_ins = time.perf_counter() - t
t = time.perf_counter()
ncon = space.connect(2.5, mesh.segments_free, block=512)
t_con = time.perf_counter() - t
kids = np.array([len(c) for c in space.children])
print("search space: %d waypoints (%.1f s), %d connections in %.1f s (children per node: mean %.1f max %d), %d mesh launches"
      % (space.n, t_ins, ncon, t_con, kids.mean(), kids.max(), mesh.L.asscpp_mesh_launch_count(mesh.h)), flush=True)
ctx = A.CoverageContext(xyz, A.default_params())
start = (3.0, -34.5, 9.0)  # coverage_heuristic_test.cpp:83
for accumulate, target in ((False, 5.0), (True, 30.0)):
    gs = SR.GpuScorer(ctx, accumulate)
    t = time.perf_counter()
    got = SR.astar_search(space, start, gs, target, 1.0, max_expansions=5000)
    t_gpu = time.perf_counter() - t
    t = time.perf_counter()
    want = SR.astar_search(space, start, OracleScorer(O.Oracle(xyz), O, accumulate), target, 1.0, max_expansions=5000)
    t_cpu = time.perf_counter() - t
    same = got is not None and want is not None and all(got[k] == want[k] for k in ("path", "coverage", "f", "expanded", "created"))
    print("accumulate=%s target %.0f%%: path of %d waypoints, coverage %.3f%%, distance %.2f m, %d expansions, %d nodes created, "
          "%d scoring launches; search %.2f s with the GPU scorer, %.2f s with the oracle scorer; identical: %s"
          % (accumulate, target, len(got["path"]), got["coverage"][-1], got["distance"], got["expanded"], got["created"],
             gs.launches, t_gpu, t_cpu, same), flush=True)
    if not same:
        sys.exit(1)
print("SEARCH OK")
