import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import asscpp_b200 as A
g = np.load(os.path.join(ROOT, "tests", "golden", "etihad_model.npz"))
sp = np.load(os.path.join(ROOT, "tests", "golden", "etihad_searchspace_2p5m.npz"))
ctx = A.CoverageContext(g["xyz"], A.default_params())
s = ctx.new_set(); s.or_views(sp["cam0"][600:620], 5)
for n in (2, 40, 400, 2450):
    poses = np.stack([sp["cam0"][:n // 2], sp["cam1"][:n // 2]], 1).reshape(-1, 7)
    for _ in range(20): ctx.eval_groups(poses, 2, covered=s)
    t = time.perf_counter()
    reps = 300
    for _ in range(reps): ctx.eval_groups(poses, 2, covered=s)
    dt = (time.perf_counter() - t) / reps
    print("views %5d: %.1f us per call, %.0f VE/s" % (n, dt * 1e6, n / dt))
