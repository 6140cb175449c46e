"""ctypes binding of oracle/_ref/libasscpp_ref.so: the REFERENCE's own sources (cscpp/src, sspp/src) compiled from
/root/reference against the stand-in third-party headers under compat/ (oracle/Makefile, oracle/ref_facade.cpp).

TEST INFRASTRUCTURE ONLY: used by tests/, tools/ and bench.py's cpu_baseline / --impl reference legs to pin the
oracle and to time the reference's CPU implementation.  Never imported by the product.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libasscpp_ref.so")
REFERENCE_ROOT = os.environ.get("ASSCPP_REFERENCE_ROOT", "/root/reference")


def available():
    """The library exists (built here where /root/reference is mounted; shipped prebuilt to the GPU box)."""
    return os.path.exists(LIB_PATH) or os.path.exists(os.path.join(REFERENCE_ROOT, "cscpp", "src", "occlusion_culling.cpp"))


def build(lib_path=LIB_PATH, defs="", objdir=None):
    """Compile the reference's sources where they lie (never copied) with the committed Makefile."""
    if os.path.exists(os.path.join(REFERENCE_ROOT, "cscpp", "src", "occlusion_culling.cpp")):
        args = ["make", "-C", _HERE, "-j8", "ref", "REF=" + REFERENCE_ROOT, "REF_OUT=" + os.path.relpath(lib_path, _HERE)]
        if defs:
            args.append("REF_DEFS=" + defs)
        if objdir:
            args.append("REF_OBJDIR=" + os.path.relpath(objdir, _HERE))
        subprocess.check_call(args, stdout=subprocess.DEVNULL)
    if not os.path.exists(lib_path):
        raise RuntimeError("oracle/_ref is not built and /root/reference is not here to build it from")
    return lib_path


_libs = {}


def lib(path=LIB_PATH):
    if path in _libs:
        return _libs[path]
    if path == LIB_PATH:
        build()
    L = C.CDLL(path)
    vp, i64, f64p, f32p, i32p, i64p = (C.c_void_p, C.c_int64, C.POINTER(C.c_double), C.POINTER(C.c_float),
                                       C.POINTER(C.c_int32), C.POINTER(C.c_int64))
    L.ref_oc_create_pcd.argtypes = [C.c_char_p]; L.ref_oc_create_pcd.restype = vp
    L.ref_oc_create_points.argtypes = [f32p, i64, C.c_float]; L.ref_oc_create_points.restype = vp
    L.ref_oc_destroy.argtypes = [vp]; L.ref_oc_destroy.restype = None
    L.ref_oc_num_points.argtypes = [vp]; L.ref_oc_num_points.restype = i64
    L.ref_oc_original_voxels.argtypes = [vp]; L.ref_oc_original_voxels.restype = C.c_float
    L.ref_oc_set_origin.argtypes = [vp, f32p]; L.ref_oc_set_origin.restype = None
    L.ref_oc_extract.argtypes = [vp, f64p, i32p, i32p, i64, i64p]; L.ref_oc_extract.restype = C.c_int
    L.ref_oc_coverage_percent_idx.argtypes = [vp, i32p, i64]; L.ref_oc_coverage_percent_idx.restype = C.c_float
    L.ref_oc_avg_accuracy_idx.argtypes = [vp, f64p, i32p, i64, f64p]; L.ref_oc_avg_accuracy_idx.restype = None
    L.ref_oc_viewport_transform.argtypes = [vp, f64p, f32p, i64, f32p]; L.ref_oc_viewport_transform.restype = None
    L.ref_oc_eval_batch.argtypes = [C.POINTER(vp), C.c_int, f64p, i64, i64p]; L.ref_oc_eval_batch.restype = C.c_int
    u64p = C.POINTER(C.c_uint64)
    L.ref_oc_create_many.argtypes = [f32p, i64, C.c_float, C.c_int, C.POINTER(vp)]; L.ref_oc_create_many.restype = None
    L.ref_oc_eval_batch_gain.argtypes = [C.POINTER(vp), C.c_int, f64p, i64, u64p, u64p, i32p, i64, i64p]
    L.ref_oc_eval_batch_gain.restype = C.c_int
    L.ref_oc_accumulate.argtypes = [vp, f64p, u64p, u64p, i32p]; L.ref_oc_accumulate.restype = C.c_int
    L.ref_robot2sensor.argtypes = [f64p, f64p, f64p, f64p]; L.ref_robot2sensor.restype = None
    L.ref_planner_create.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_double, C.c_double, C.c_double]
    L.ref_planner_create.restype = vp
    L.ref_planner_destroy.argtypes = [vp]; L.ref_planner_destroy.restype = None
    L.ref_planner_dynamic_sampling.argtypes = [vp, f64p, f64p, C.c_double, C.c_double]; L.ref_planner_dynamic_sampling.restype = None
    L.ref_planner_regular_grid.argtypes = [vp, f64p, f64p, C.c_double]; L.ref_planner_regular_grid.restype = None
    L.ref_planner_load_grid.argtypes = [vp, C.c_char_p, C.c_char_p, C.c_char_p]; L.ref_planner_load_grid.restype = None
    L.ref_planner_search_space.argtypes = [vp, f64p, f64p, f64p, i64]; L.ref_planner_search_space.restype = i64
    L.ref_planner_num_connections.argtypes = [vp]; L.ref_planner_num_connections.restype = i64
    L.ref_planner_point_cloud_diff.argtypes = [vp, f32p, i64, f32p, i64, i64p]; L.ref_planner_point_cloud_diff.restype = C.c_double
    L.ref_planner_clusters.argtypes = [vp, f32p, i64, i32p, f64p, i64]; L.ref_planner_clusters.restype = i64
    L.ref_planner_set_accuracy_range.argtypes = [vp]; L.ref_planner_set_accuracy_range.restype = None
    L.ref_planner_search.argtypes = [vp, f64p, f64p, i64]; L.ref_planner_search.restype = i64
    L.ref_planner_infogain_chain.argtypes = [vp, f64p, f64p, C.c_int, i64, f64p]; L.ref_planner_infogain_chain.restype = i64
    _libs[path] = L
    return L


def _p(a, ty):
    return None if a is None else a.ctypes.data_as(C.POINTER(ty))


class RefOcclusionCulling:
    """The reference's OcclusionCulling object (cscpp/src/occlusion_culling.cpp), one per host thread."""

    def __init__(self, xyz=None, pcd_path=None, leaf=0.0, nthreads=1, origin=None, lib_path=LIB_PATH):
        self.L = lib(lib_path)
        self.handles = []
        if pcd_path is None and nthreads > 1:
            a = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
            arr = (C.c_void_p * nthreads)()
            self.L.ref_oc_create_many(_p(a, C.c_float), len(a), float(leaf), nthreads, arr)
            self.handles = [arr[i] for i in range(nthreads)]
            nthreads = 0
        for _ in range(nthreads if self.handles else max(1, nthreads)):
            if pcd_path is not None:
                h = self.L.ref_oc_create_pcd(pcd_path.encode())
            else:
                a = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
                h = self.L.ref_oc_create_points(_p(a, C.c_float), len(a), float(leaf))
            if origin is not None:
                o = np.ascontiguousarray(origin, np.float32)
                self.L.ref_oc_set_origin(h, _p(o, C.c_float))
            self.handles.append(h)
        self.n = self.L.ref_oc_num_points(self.handles[0])
        self.original_voxels = self.L.ref_oc_original_voxels(self.handles[0])

    def close(self):
        for h in self.handles:
            self.L.ref_oc_destroy(h)
        self.handles = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def extract(self, pose):
        """(survivor indices, visible indices) of extractVisibleSurface for one pose."""
        pose = np.ascontiguousarray(pose, np.float64)
        surv = np.zeros(self.n, np.int32)
        vis = np.zeros(self.n, np.int32)
        cnt = np.zeros(2, np.int64)
        rc = self.L.ref_oc_extract(self.handles[0], _p(pose, C.c_double), _p(surv, C.c_int32), _p(vis, C.c_int32), self.n,
                                   _p(cnt, C.c_int64))
        if rc:
            raise RuntimeError("reference cloud is not a subsequence of the model cloud")
        return surv[:cnt[0]].copy(), vis[:cnt[1]].copy()

    def eval_batch(self, poses):
        """stats [n, 3]: survivors, visible, hash of the visible index set (same mix64 as the oracle / the GPU)."""
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 7)
        st = np.zeros((len(poses), 3), np.int64)
        arr = (C.c_void_p * len(self.handles))(*self.handles)
        rc = self.L.ref_oc_eval_batch(arr, len(self.handles), _p(poses, C.c_double), len(poses), _p(st, C.c_int64))
        if rc:
            raise RuntimeError("ref_oc_eval_batch failed")
        return st

    def eval_batch_gain(self, poses, cov_pts, cov_vox, pt_cov, n_cov):
        """stats [n, 5]: survivors, visible, hash, gain_pts, gain_vox against the covered bitsets (uint64 words)."""
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 7)
        st = np.zeros((len(poses), 5), np.int64)
        arr = (C.c_void_p * len(self.handles))(*self.handles)
        rc = self.L.ref_oc_eval_batch_gain(arr, len(self.handles), _p(poses, C.c_double), len(poses), _p(cov_pts, C.c_uint64),
                                           _p(cov_vox, C.c_uint64), _p(pt_cov, C.c_int32), int(n_cov), _p(st, C.c_int64))
        if rc:
            raise RuntimeError("ref_oc_eval_batch_gain failed")
        return st

    def accumulate(self, pose, cov_pts, cov_vox, pt_cov):
        pose = np.ascontiguousarray(pose, np.float64)
        if self.L.ref_oc_accumulate(self.handles[0], _p(pose, C.c_double), _p(cov_pts, C.c_uint64), _p(cov_vox, C.c_uint64),
                                    _p(pt_cov, C.c_int32)):
            raise RuntimeError("ref_oc_accumulate failed")

    def coverage_percent_idx(self, idx):
        idx = np.ascontiguousarray(idx, np.int32)
        return float(self.L.ref_oc_coverage_percent_idx(self.handles[0], _p(idx, C.c_int32), len(idx)))

    def avg_accuracy_idx(self, pose, idx):
        pose = np.ascontiguousarray(pose, np.float64)
        idx = np.ascontiguousarray(idx, np.int32)
        out = np.zeros(2, np.float64)
        self.L.ref_oc_avg_accuracy_idx(self.handles[0], _p(pose, C.c_double), _p(idx, C.c_int32), len(idx), _p(out, C.c_double))
        return out[0], out[1]

    def viewport_transform(self, pose, xyz):
        pose = np.ascontiguousarray(pose, np.float64)
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros_like(xyz)
        self.L.ref_oc_viewport_transform(self.handles[0], _p(pose, C.c_double), _p(xyz, C.c_float), len(xyz), _p(out, C.c_float))
        return out


def robot2sensor(robot_pose, offset, rpy, lib_path=LIB_PATH):
    rp = np.ascontiguousarray(robot_pose, np.float64)
    off = np.ascontiguousarray(offset, np.float64)
    r = np.ascontiguousarray(rpy, np.float64)
    out = np.zeros(7, np.float64)
    lib(lib_path).ref_robot2sensor(_p(rp, C.c_double), _p(off, C.c_double), _p(r, C.c_double), _p(out, C.c_double))
    return out


class RefPlanner:
    """The reference's PathPlanner + CoveragePathPlanningHeuristic wired as its demo does
    (cscpp/src/coverage_heuristic_test.cpp:73-137)."""

    def __init__(self, obj_path, pcd_path, heuristic_type=0, target=5.0, tolerance=1.0, conn_radius=2.5, lib_path=LIB_PATH):
        self.L = lib(lib_path)
        self.h = self.L.ref_planner_create(obj_path.encode(), pcd_path.encode(), heuristic_type, target, tolerance, conn_radius)

    def close(self):
        if self.h:
            self.L.ref_planner_destroy(self.h)
            self.h = None

    def dynamic_sampling(self, grid_start=(-18.0, -25.0, 1.0), grid_size=(36.0, 50.0, 15.0), start_res=4.5, decrement=1.5):
        gs = np.ascontiguousarray(grid_start, np.float64)
        sz = np.ascontiguousarray(grid_size, np.float64)
        self.L.ref_planner_dynamic_sampling(self.h, _p(gs, C.c_double), _p(sz, C.c_double), start_res, decrement)

    def regular_grid(self, grid_start, grid_size, res):
        gs = np.ascontiguousarray(grid_start, np.float64)
        sz = np.ascontiguousarray(grid_size, np.float64)
        self.L.ref_planner_regular_grid(self.h, _p(gs, C.c_double), _p(sz, C.c_double), res)

    def load_grid(self, robot_file, sensor0_file, sensor1_file):
        self.L.ref_planner_load_grid(self.h, robot_file.encode(), sensor0_file.encode(), sensor1_file.encode())

    def search_space(self):
        n = self.L.ref_planner_search_space(self.h, None, None, None, 0)
        rows = np.zeros((n, 4), np.float64)
        robot = np.zeros((n, 7), np.float64)
        sens = np.zeros((n, 2, 7), np.float64)
        self.L.ref_planner_search_space(self.h, _p(rows, C.c_double), _p(robot, C.c_double), _p(sens, C.c_double), n)
        return rows, robot, sens

    def point_cloud_diff(self, xyz):
        """pointCloudDiff of a covered cloud: (uncovered %, cell centres of the uncovered model voxels)."""
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros((200000, 3), np.float32)
        n = C.c_int64()
        pct = self.L.ref_planner_point_cloud_diff(self.h, _p(xyz, C.c_float), len(xyz), _p(out, C.c_float), len(out), C.byref(n))
        return pct, out[:n.value].copy()

    def clusters(self, xyz):
        """clusteringPointCloud + findClusterBB: (labels per point, boxes [k, 6])."""
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        labels = np.zeros(len(xyz), np.int32)
        boxes = np.zeros((4096, 6), np.float64)
        k = self.L.ref_planner_clusters(self.h, _p(xyz, C.c_float), len(xyz), _p(labels, C.c_int32), _p(boxes, C.c_double), len(boxes))
        return labels, boxes[:k].copy()

    def num_connections(self):
        return self.L.ref_planner_num_connections(self.h)

    def set_accuracy_range(self):
        self.L.ref_planner_set_accuracy_range(self.h)

    def search(self, start=(3.0, -34.5, 9.0)):
        st = np.ascontiguousarray(start, np.float64)
        rows = np.zeros((100000, 10), np.float64)
        n = self.L.ref_planner_search(self.h, _p(st, C.c_double), _p(rows, C.c_double), len(rows))
        return rows[:n].copy()

    def infogain_chain(self, robot_poses, sensor_poses):
        """InfoGainVolumetricH (heuristic.cpp:643-759, 1128-1225) of a chain of nodes, node i's parent = the real node
        i-1 (the form in which the reference's block is well defined).  sensor_poses [k, ns, 7].
        Rows {totalEntroby, coveredVolume, coverage, f_value, distance, coveredVoxelsNum}."""
        rp = np.ascontiguousarray(robot_poses, np.float64).reshape(-1, 7)
        sp = np.ascontiguousarray(sensor_poses, np.float64).reshape(len(rp), -1, 7)
        out = np.zeros((len(rp), 6), np.float64)
        self.L.ref_planner_infogain_chain(self.h, _p(rp, C.c_double), _p(sp, C.c_double), sp.shape[1], len(rp), _p(out, C.c_double))
        return out
