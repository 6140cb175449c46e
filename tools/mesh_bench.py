"""Times the mesh sample filters (asscpp_mesh_filter / asscpp_mesh_segments_free) on the Etihad mesh against
the oracle on the host cores.  Usage: python tools/mesh_bench.py [n_queries]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import asscpp_b200 as A  # noqa: E402
from oracle import oracle as O  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
g = np.load(os.path.join(ROOT, "tests", "golden", "etihad_mesh.npz"))
rng = np.random.default_rng(0)
pos = np.stack([rng.uniform(-22, 22, n), rng.uniform(-25, 25, n), rng.uniform(-3, 14, n)], 1)
dst = np.stack([rng.uniform(-22, 22, n), rng.uniform(-25, 25, n), rng.uniform(-3, 14, n)], 1)
gm, om = A.Mesh(g["vertices"], g["triangles"]), O.Mesh(g["vertices"], g["triangles"])
gm.filter(pos[:64]); gm.segments_free(pos[:64], dst[:64])
t = time.perf_counter(); gh, gs, go = gm.filter(pos); t_f = time.perf_counter() - t
t = time.perf_counter(); gf = gm.segments_free(pos, dst); t_s = time.perf_counter() - t
ncpu = min(n, 2000)
cores = os.cpu_count() or 1
t = time.perf_counter(); oh, os_, oo, _ = om.filter(pos[:ncpu], nthreads=cores); c_f = time.perf_counter() - t
t = time.perf_counter(); of = om.segments_free(pos[:ncpu], dst[:ncpu], nthreads=cores); c_s = time.perf_counter() - t
assert np.array_equal(gh[:ncpu], oh) and np.array_equal(gs[:ncpu].view(np.uint64), os_.view(np.uint64))
assert np.array_equal(gf[:ncpu], of)
nt = gm.n_triangles
print("triangles %d queries %d" % (nt, n))
print("filter   gpu %.3f ms  %.3g query-triangle pairs/s | cpu(%d thr) %.3g pairs/s" % (
    t_f * 1e3, n * nt / t_f, cores, ncpu * nt / c_f))
print("segments gpu %.3f ms  %.3g pairs/s | cpu %.3g pairs/s" % (t_s * 1e3, n * nt / t_s, ncpu * nt / c_s))
