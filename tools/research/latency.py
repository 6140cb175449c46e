"""Small-batch latency of the grouped evaluation (the A* use: the sensors of the children of one expansion).
Per batch size: wall time per call through host buffers, device time on the context's stream (CUDA events around the
whole call) and the time of eval_views_kernel itself."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import asscpp_b200 as A
g = np.load(os.path.join(ROOT, "tests", "golden", "etihad_model.npz"))
sp = np.load(os.path.join(ROOT, "tests", "golden", "etihad_searchspace_2p5m.npz"))
ctx = A.CoverageContext(g["xyz"], A.default_params())
s = ctx.new_set(); s.or_views(sp["cam0"][600:620], 5)
print("split-group mode up to %s views" % os.environ.get("ASSCPP_SPLIT_GROUP_MAX", "4096 (default)"))
for n in (2, 40, 400, 2450):
    poses = np.stack([sp["cam0"][:n // 2], sp["cam1"][:n // 2]], 1).reshape(-1, 7)
    for _ in range(20): ctx.eval_groups(poses, 2, covered=s)
    t = time.perf_counter()
    reps = 300
    for _ in range(reps): ctx.eval_groups(poses, 2, covered=s)
    dt = (time.perf_counter() - t) / reps
    # the same two C-ABI calls with the arguments built once (what a C++ caller pays: no numpy / ctypes marshalling)
    import ctypes as C
    o = A.EvalOpts(s.handle, -1, 5, 2, 0, 0, 0)
    pp = np.ascontiguousarray(poses, np.float64); pv = pp.ctypes.data_as(C.c_void_p)
    out = np.zeros(n, A.RESULT_DTYPE); ov = out.ctypes.data_as(C.c_void_p)
    f_eval, f_get, h, ob, nn = ctx.L.asscpp_eval_ex, ctx.L.asscpp_get_results, ctx.h, C.byref(o), C.c_int64(n)
    for _ in range(20): f_eval(h, pv, nn, ob); f_get(h, ov, nn)
    t = time.perf_counter()
    for _ in range(reps): f_eval(h, pv, nn, ob); f_get(h, ov, nn)
    dt_abi = (time.perf_counter() - t) / reps
    dev, kern = [], []
    for _ in range(50):
        ctx.timer_begin()
        ctx.eval_ex(poses, covered=s, group_size=2)
        dev.append(ctx.timer_end())
        kern.append(ctx.eval_kernel_ms()[0])
        ctx.gains(n)
    print("views %5d: %.1f us per call, %.0f VE/s; C-ABI calls alone %.1f us; stream %.1f us, eval_views_kernel %.1f us" % (
        n, dt * 1e6, n / dt, dt_abi * 1e6, 1e3 * float(np.median(dev)), 1e3 * float(np.median(kern))))
