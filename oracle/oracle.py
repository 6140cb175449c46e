"""ctypes binding of the CPU oracle (oracle/asscpp_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
PARITY UNPINNED (see asscpp_oracle.h).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libasscpp_oracle.so")

ST_SURVIVORS, ST_VISIBLE, ST_CARRY, ST_UNIQUE, ST_CELLS, ST_STEPS, ST_GAIN_PTS, ST_GAIN_VOX, ST_HASH, ST_SURV_HASH = range(10)
ST_COUNT = 10
EPS_COUNT = 5


class OrcParams(C.Structure):
    _fields_ = [
        ("leaf", C.c_float),
        ("hfov_deg", C.c_float),
        ("vfov_deg", C.c_float),
        ("near_dist", C.c_float),
        ("far_dist", C.c_float),
        ("pcd_origin", C.c_float * 3),
        ("cov_leaf", C.c_float),
    ]


def make_params(leaf=0.5, hfov=58.0, vfov=45.0, near=0.7, far=6.0, pcd_origin=(0.0, 0.0, 0.0), cov_leaf=0.5):
    p = OrcParams()
    p.leaf, p.hfov_deg, p.vfov_deg, p.near_dist, p.far_dist = leaf, hfov, vfov, near, far
    p.pcd_origin[:] = pcd_origin
    p.cov_leaf = cov_leaf
    return p


def build(force=False):
    """Compile the oracle with the committed Makefile (g++ only, no reference sources)."""
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(
        os.path.join(_HERE, "asscpp_oracle.cpp")
    ):
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, i64, i32p, i64p, u64p, f32p, f64p = (
            C.c_void_p, C.c_int64, C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_uint64),
            C.POINTER(C.c_float), C.POINTER(C.c_double))
        L.orc_create.argtypes = [f32p, i64, C.POINTER(OrcParams), C.POINTER(vp)]
        L.orc_create.restype = C.c_int
        L.orc_destroy.argtypes = [vp]
        L.orc_destroy.restype = None
        for name in ("orc_num_points", "orc_num_cov_voxels", "orc_num_occ_voxels"):
            getattr(L, name).argtypes = [vp]
            getattr(L, name).restype = i64
        L.orc_model_box.argtypes = [vp, i32p, i32p]
        L.orc_model_box.restype = None
        L.orc_point_cov_ids.argtypes = [vp, i32p]
        L.orc_point_cov_ids.restype = None
        L.orc_planes.argtypes = [C.POINTER(OrcParams), f64p, f32p, f32p]
        L.orc_planes.restype = None
        L.orc_corners.argtypes = [C.POINTER(OrcParams), f64p, f32p]
        L.orc_corners.restype = None
        L.orc_eval_view.argtypes = [vp, f64p, C.c_int, u64p, u64p, i32p, i32p, i64p, i64p]
        L.orc_eval_view.restype = C.c_int
        L.orc_eval_batch.argtypes = [vp, f64p, i64, C.c_int, C.c_int, u64p, u64p, i64p]
        L.orc_eval_batch.restype = C.c_int
        L.orc_accumulate.argtypes = [vp, f64p, i64, C.c_int, i64, u64p, u64p]
        L.orc_accumulate.restype = C.c_int
        L.orc_coverage_percent_idx.argtypes = [vp, i32p, i64]
        L.orc_coverage_percent_idx.restype = C.c_float
        L.orc_coverage_percent_vox.argtypes = [vp, u64p]
        L.orc_coverage_percent_vox.restype = C.c_float
        L.orc_popcount.argtypes = [u64p, i64]
        L.orc_popcount.restype = i64
        L.orc_robot2sensor.argtypes = [f64p, f64p, f64p, f64p]
        L.orc_robot2sensor.restype = None
        L.orc_pose_rpy.argtypes = [f64p, f64p]
        L.orc_pose_rpy.restype = None
        L.orc_viewport_transform.argtypes = [f64p, f32p, i64, f32p]
        L.orc_viewport_transform.restype = None
        L.orc_accuracy_idx.argtypes = [vp, f64p, i32p, i64, f64p, f64p, f32p, f32p]
        L.orc_accuracy_idx.restype = None
        L.orc_uncovered.argtypes = [vp, u64p, i32p, f32p]
        L.orc_uncovered.restype = i64
        L.orc_uncovered_ex.argtypes = [vp, u64p, C.c_int, i32p, f32p]
        L.orc_uncovered_ex.restype = i64
        L.orc_euclidean_clusters.argtypes = [f32p, i64, C.c_double, C.c_int32, C.c_int32, i32p, i64p]
        L.orc_euclidean_clusters.restype = C.c_int32
        L.orc_cluster_bb.argtypes = [f32p, i64, C.c_float, f64p, f64p]
        L.orc_cluster_bb.restype = None
        u8p = C.POINTER(C.c_uint8)
        L.orc_mesh_create.argtypes = [f32p, i64, i32p, i64]
        L.orc_mesh_create.restype = vp
        L.orc_mesh_destroy.argtypes = [vp]
        L.orc_mesh_destroy.restype = None
        L.orc_mesh_num_triangles.argtypes = [vp]
        L.orc_mesh_num_triangles.restype = i64
        L.orc_mesh_filter.argtypes = [vp, f64p, i64, f64p, C.c_double, C.c_double, C.c_int, i32p, f64p, u8p, f64p]
        L.orc_mesh_filter.restype = None
        L.orc_mesh_segments_free.argtypes = [vp, f64p, f64p, i64, C.c_int, u8p]
        L.orc_mesh_segments_free.restype = None
        L.orc_mix64.argtypes = [C.c_uint64]
        L.orc_mix64.restype = C.c_uint64
        _lib = L
    return _lib


def _p(a, ty):
    return None if a is None else a.ctypes.data_as(C.POINTER(ty))


class Oracle:
    """One model + parameter set, evaluated on the host."""

    def __init__(self, xyz, params=None):
        self.L = lib()
        self.xyz = np.ascontiguousarray(xyz, dtype=np.float32).reshape(-1, 3)
        self.params = params or make_params()
        self.h = C.c_void_p()
        rc = self.L.orc_create(_p(self.xyz, C.c_float), self.xyz.shape[0], C.byref(self.params), C.byref(self.h))
        if rc != 0:
            raise RuntimeError("orc_create failed: %d" % rc)
        self.n = self.xyz.shape[0]
        self.n_cov = self.L.orc_num_cov_voxels(self.h)
        self.n_occ = self.L.orc_num_occ_voxels(self.h)

    def __del__(self):
        if getattr(self, "h", None) and self.h.value:
            self.L.orc_destroy(self.h)
            self.h = C.c_void_p()

    def pts_words(self):
        return (self.n + 63) // 64

    def vox_words(self):
        return (self.n_cov + 63) // 64

    def planes(self, pose):
        pose = np.ascontiguousarray(pose, dtype=np.float64)
        pl = np.zeros(24, np.float32)
        T = np.zeros(3, np.float32)
        self.L.orc_planes(C.byref(self.params), _p(pose, C.c_double), _p(pl, C.c_float), _p(T, C.c_float))
        return pl.reshape(6, 4), T

    def corners(self, pose):
        pose = np.ascontiguousarray(pose, dtype=np.float64)
        cs = np.zeros(24, np.float32)
        self.L.orc_corners(C.byref(self.params), _p(pose, C.c_double), _p(cs, C.c_float))
        return cs.reshape(8, 3)

    def cov_ids(self):
        out = np.zeros(self.n, np.int32)
        self.L.orc_point_cov_ids(self.h, _p(out, C.c_int32))
        return out

    def eval_view(self, pose, memo=False, covered_pts=None, covered_vox=None, want_eps=False):
        pose = np.ascontiguousarray(pose, dtype=np.float64)
        surv = np.zeros(self.n, np.int32)
        vis = np.zeros(self.n, np.int32)
        st = np.zeros(ST_COUNT, np.int64)
        eps = np.zeros(EPS_COUNT, np.int64) if want_eps else None
        rc = self.L.orc_eval_view(self.h, _p(pose, C.c_double), int(memo), _p(covered_pts, C.c_uint64),
                                  _p(covered_vox, C.c_uint64), _p(surv, C.c_int32), _p(vis, C.c_int32),
                                  _p(st, C.c_int64), _p(eps, C.c_int64))
        if rc != 0:
            raise RuntimeError("orc_eval_view failed: %d" % rc)
        out = {"survivors": surv[: st[ST_SURVIVORS]].copy(), "visible": vis[: st[ST_VISIBLE]].copy(), "stats": st}
        if want_eps:
            out["eps"] = eps
        return out

    def eval_batch(self, poses, memo=True, nthreads=1, covered_pts=None, covered_vox=None):
        poses = np.ascontiguousarray(poses, dtype=np.float64).reshape(-1, 7)
        st = np.zeros((poses.shape[0], ST_COUNT), np.int64)
        rc = self.L.orc_eval_batch(self.h, _p(poses, C.c_double), poses.shape[0], int(memo), int(nthreads),
                                   _p(covered_pts, C.c_uint64), _p(covered_vox, C.c_uint64), _p(st, C.c_int64))
        if rc != 0:
            raise RuntimeError("orc_eval_batch failed: %d" % rc)
        return st

    def accumulate(self, poses, covered_pts, covered_vox, min_visible=5, nthreads=1):
        poses = np.ascontiguousarray(poses, dtype=np.float64).reshape(-1, 7)
        rc = self.L.orc_accumulate(self.h, _p(poses, C.c_double), poses.shape[0], int(nthreads), int(min_visible),
                                   _p(covered_pts, C.c_uint64), _p(covered_vox, C.c_uint64))
        if rc != 0:
            raise RuntimeError("orc_accumulate failed: %d" % rc)

    def coverage_percent_idx(self, idx):
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        return float(self.L.orc_coverage_percent_idx(self.h, _p(idx, C.c_int32), idx.shape[0]))

    def uncovered(self, covered_vox=None, pcl_aliasing=False):
        """pointCloudDiff (heuristic.cpp:88-145): (ids, centres) of the uncovered coverage voxels."""
        ids = np.zeros(max(self.n_cov, 1), np.int32)
        ctr = np.zeros((max(self.n_cov, 1), 3), np.float32)
        n = lib().orc_uncovered_ex(self.h, _p(covered_vox, C.c_uint64), int(pcl_aliasing), _p(ids, C.c_int32), _p(ctr, C.c_float))
        return ids[:n].copy(), ctr[:n].copy()

    def accuracy_idx(self, pose, idx):
        """(err_sum_depth, err_sum_z, depth_min, depth_max) over the points idx in that order:
        calcAvgAccuracy(cloud, pose) = err_sum_depth / n, calcAvgAccuracy(cloud) = err_sum_z / n."""
        pose = np.ascontiguousarray(pose, np.float64)
        idx = np.ascontiguousarray(idx, np.int32)
        sd, sz = C.c_double(0), C.c_double(0)
        lo, hi = C.c_float(0), C.c_float(0)
        lib().orc_accuracy_idx(self.h, _p(pose, C.c_double), _p(idx, C.c_int32), len(idx), C.byref(sd), C.byref(sz),
                               C.byref(lo), C.byref(hi))
        return sd.value, sz.value, lo.value, hi.value

    def coverage_percent_vox(self, covered_vox):
        return float(self.L.orc_coverage_percent_vox(self.h, _p(covered_vox, C.c_uint64)))


def pose_rpy(pose):
    """tf::Matrix3x3(q).getRPY as used by occlusion_culling.cpp:497-500."""
    pose = np.ascontiguousarray(pose, np.float64)
    out = np.zeros(3, np.float64)
    lib().orc_pose_rpy(_p(pose, C.c_double), _p(out, C.c_double))
    return out


def viewport_transform(pose, xyz):
    """pointCloudViewportTransform, occlusion_culling.cpp:538-618."""
    pose = np.ascontiguousarray(pose, np.float64)
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    out = np.zeros_like(xyz)
    lib().orc_viewport_transform(_p(pose, C.c_double), _p(xyz, C.c_float), len(xyz), _p(out, C.c_float))
    return out


class Mesh:
    """Mesh filters of the adaptive loop (heuristic.cpp:300-335), CPU restatement."""

    def __init__(self, vertices, triangles):
        v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
        t = np.ascontiguousarray(triangles, np.int32).reshape(-1, 3)
        self.h = lib().orc_mesh_create(_p(v, C.c_float), len(v), _p(t, C.c_int32), len(t))
        if not self.h:
            raise ValueError("bad mesh")
        self.n_triangles = lib().orc_mesh_num_triangles(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_mesh_destroy(self.h)
            self.h = None

    def filter(self, positions, through=(100.0, 10.0, 56.0), min_dist=1.0, max_dist=4.0, nthreads=1):
        pos = np.ascontiguousarray(positions, np.float64).reshape(-1, 3)
        thr = np.ascontiguousarray(through, np.float64)
        hits = np.zeros(len(pos), np.int32)
        sqd = np.zeros(len(pos), np.float64)
        ok = np.zeros(len(pos), np.uint8)
        graze = np.zeros(len(pos), np.float64)
        lib().orc_mesh_filter(self.h, _p(pos, C.c_double), len(pos), _p(thr, C.c_double), min_dist, max_dist, nthreads,
                              _p(hits, C.c_int32), _p(sqd, C.c_double), _p(ok, C.c_uint8), _p(graze, C.c_double))
        return hits, sqd, ok.astype(bool), graze

    def segments_free(self, a, b, nthreads=1):
        a = np.ascontiguousarray(a, np.float64).reshape(-1, 3)
        b = np.ascontiguousarray(b, np.float64).reshape(-1, 3)
        out = np.zeros(len(a), np.uint8)
        lib().orc_mesh_segments_free(self.h, _p(a, C.c_double), _p(b, C.c_double), len(a), nthreads, _p(out, C.c_uint8))
        return out.astype(bool)


def euclidean_clusters(xyz, tol=0.8, min_size=5, max_size=50000):
    """clusteringPointCloud (heuristic.cpp:147-190): labels (cluster rank, largest first, or -1), count, exact ties."""
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    labels = np.zeros(len(xyz), np.int32)
    ties = C.c_int64(0)
    k = lib().orc_euclidean_clusters(_p(xyz, C.c_float), len(xyz), tol, min_size, max_size, _p(labels, C.c_int32),
                                     C.byref(ties))
    return labels, int(k), ties.value


def cluster_bb(xyz, leaf=0.5):
    """findClusterBB (heuristic.cpp:218-246): (grid_start[3], grid_size[3])."""
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    start, size = np.zeros(3), np.zeros(3)
    lib().orc_cluster_bb(_p(xyz, C.c_float), len(xyz), leaf, _p(start, C.c_double), _p(size, C.c_double))
    return start, size


def robot2sensor(robot_pose, offset, rpy):
    L = lib()
    rp = np.ascontiguousarray(robot_pose, dtype=np.float64)
    off = np.ascontiguousarray(offset, dtype=np.float64)
    r = np.ascontiguousarray(rpy, dtype=np.float64)
    out = np.zeros(7, np.float64)
    L.orc_robot2sensor(_p(rp, C.c_double), _p(off, C.c_double), _p(r, C.c_double), _p(out, C.c_double))
    return out


def set_hash(idx):
    """Order-independent hash used by both sides: sum of mix64(idx) mod 2^64."""
    x = np.asarray(idx, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        x = x ^ (x >> np.uint64(31))
        return int(np.sum(x, dtype=np.uint64))
