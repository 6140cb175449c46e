"""asscpp_b200 -- B200-native viewpoint coverage evaluation (frustum cull -> 3D-DDA occlusion ->
coverage gain), the data-parallel hot path of kucars/asscpp.

The product is the C-ABI library `asscpp_b200/lib/libasscpp_b200.so` (include/asscpp_gpu.h) and the
C++ shim of the reference's OcclusionCulling class (include/cscpp/occlusion_culling.h).  This Python
package is a thin ctypes mirror used by the tests, the benchmark and Python callers; it adds no
compute of its own and has no fallback when the CUDA library is absent.
"""
import ctypes as C

import numpy as np

from . import _abi
from ._abi import (ACCURACY_DTYPE, CLUSTER_DTYPE, EVAL_ACCURACY, EVAL_BRUTE, EVAL_FORCE_GLOBAL, EVAL_HASH, EVAL_VIEW_CELLS, RESULT_DTYPE,
                   VS_GRID_GLOBAL, VS_GRID_SHARED, VS_ORDERED, EvalOpts, ModelInfo, Params)

__all__ = [
    "CoverageContext", "CoveredSet", "Mesh", "AsscppError", "default_params", "synth_structure", "synth_views",
    "robot2sensor", "frustum_planes", "pose_rpy", "viewport_transform", "EVAL_HASH", "EVAL_BRUTE",
    "EVAL_FORCE_GLOBAL", "EVAL_ACCURACY", "EVAL_VIEW_CELLS", "RESULT_DTYPE", "ACCURACY_DTYPE",
]

SEED_STRUCTURE, SEED_POINTS, SEED_VIEWS = 0xA55C0001, 0xA55C0002, 0xA55C0003


class AsscppError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("asscpp error %d: %s" % (code, msg))
        self.code = code


def _ptr(a, ty):
    return None if a is None else a.ctypes.data_as(C.POINTER(ty))


def default_params(**kw):
    """Reference literals (occlusion_culling.cpp:41,73-76; heuristic.cpp:479), overridable."""
    p = Params()
    _abi.load().asscpp_default_params(C.byref(p))
    for k, v in kw.items():
        if k == "pcd_origin":
            p.pcd_origin[:] = v
        else:
            setattr(p, k, v)
    return p


def synth_structure(n, structure_seed=SEED_STRUCTURE, point_seed=SEED_POINTS):
    out = np.empty((n, 3), np.float32)
    rc = _abi.load().asscpp_synth_structure_v1(structure_seed, point_seed, n, _ptr(out, C.c_float))
    if rc:
        raise AsscppError(rc, "synth_structure")
    return out


def synth_views(nviews, structure_seed=SEED_STRUCTURE, view_seed=SEED_VIEWS):
    out = np.empty((nviews, 7), np.float64)
    rc = _abi.load().asscpp_synth_views_v1(structure_seed, view_seed, nviews, _ptr(out, C.c_double))
    if rc:
        raise AsscppError(rc, "synth_views")
    return out


def robot2sensor(robot_pose, offset, rpy):
    """sspp/src/sensors.cpp:47-90."""
    rp = np.ascontiguousarray(robot_pose, np.float64)
    off = np.ascontiguousarray(offset, np.float64)
    r = np.ascontiguousarray(rpy, np.float64)
    out = np.zeros(7, np.float64)
    rc = _abi.load().asscpp_robot2sensor(_ptr(rp, C.c_double), _ptr(off, C.c_double), _ptr(r, C.c_double),
                                         _ptr(out, C.c_double))
    if rc:
        raise AsscppError(rc, "robot2sensor")
    return out


def robot2sensor_batch(robot_poses, offset, rpy):
    """sspp/src/sensors.cpp:47-90 for a whole search space in one call ([n, 7] -> [n, 7])."""
    rp = np.ascontiguousarray(robot_poses, np.float64).reshape(-1, 7)
    off = np.ascontiguousarray(offset, np.float64)
    r = np.ascontiguousarray(rpy, np.float64)
    out = np.zeros_like(rp)
    rc = _abi.load().asscpp_robot2sensor_batch(_ptr(rp, C.c_double), len(rp), _ptr(off, C.c_double),
                                               _ptr(r, C.c_double), _ptr(out, C.c_double))
    if rc:
        raise AsscppError(rc, "robot2sensor_batch")
    return out


def frustum_planes(params, pose):
    pose = np.ascontiguousarray(pose, np.float64)
    pl = np.zeros(24, np.float32)
    o = np.zeros(3, np.float32)
    rc = _abi.load().asscpp_frustum_planes(C.byref(params), _ptr(pose, C.c_double), _ptr(pl, C.c_float),
                                           _ptr(o, C.c_float))
    if rc:
        raise AsscppError(rc, "frustum_planes")
    return pl.reshape(6, 4), o


def pose_rpy(pose):
    """tf::Matrix3x3(q).getRPY of a pose (host)."""
    pose = np.ascontiguousarray(pose, np.float64)
    out = np.zeros(3, np.float64)
    rc = _abi.load().asscpp_pose_rpy(_ptr(pose, C.c_double), _ptr(out, C.c_double))
    if rc:
        raise AsscppError(rc, "pose_rpy")
    return out


def viewport_transform(pose, xyz):
    """OcclusionCulling::pointCloudViewportTransform (host)."""
    pose = np.ascontiguousarray(pose, np.float64)
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    out = np.zeros_like(xyz)
    rc = _abi.load().asscpp_viewport_transform(_ptr(pose, C.c_double), _ptr(xyz, C.c_float), len(xyz),
                                               _ptr(out, C.c_float))
    if rc:
        raise AsscppError(rc, "viewport_transform")
    return out


class Mesh:
    """Triangle mesh of the structure on the device: the CGAL sample filters of the adaptive loop
    (coverage_path_planning_heuristic.cpp:300-335)."""

    def __init__(self, vertices, triangles, device=0):
        self.L = _abi.load()
        v = np.ascontiguousarray(vertices, np.float32).reshape(-1, 3)
        t = np.ascontiguousarray(triangles, np.int32).reshape(-1, 3)
        h = C.c_void_p()
        rc = self.L.asscpp_mesh_create(device, _ptr(v, C.c_float), len(v), _ptr(t, C.c_int32), len(t), C.byref(h))
        if rc:
            raise AsscppError(rc, (self.L.asscpp_mesh_last_error(None) or b"").decode())
        self.h = h
        self.n_triangles = self.L.asscpp_mesh_num_triangles(self.h)

    def close(self):
        if getattr(self, "h", None):
            self.L.asscpp_mesh_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc:
            raise AsscppError(rc, (self.L.asscpp_mesh_last_error(self.h) or b"").decode())

    def filter(self, positions, through=(100.0, 10.0, 56.0), min_dist=1.0, max_dist=4.0):
        """(crossings of the inside-test ray, squared distance to the mesh, pass) per position."""
        pos = np.ascontiguousarray(positions, np.float64).reshape(-1, 3)
        thr = np.ascontiguousarray(through, np.float64)
        hits = np.zeros(len(pos), np.int32)
        sqd = np.zeros(len(pos), np.float64)
        ok = np.zeros(len(pos), np.uint8)
        self._ck(self.L.asscpp_mesh_filter(self.h, _ptr(pos, C.c_double), len(pos), _ptr(thr, C.c_double), min_dist,
                                           max_dist, _ptr(hits, C.c_int32), _ptr(sqd, C.c_double), _ptr(ok, C.c_uint8)))
        return hits, sqd, ok.astype(bool)

    def segments_free(self, a, b):
        a = np.ascontiguousarray(a, np.float64).reshape(-1, 3)
        b = np.ascontiguousarray(b, np.float64).reshape(-1, 3)
        out = np.zeros(len(a), np.uint8)
        self._ck(self.L.asscpp_mesh_segments_free(self.h, _ptr(a, C.c_double), _ptr(b, C.c_double), len(a),
                                                  _ptr(out, C.c_uint8)))
        return out.astype(bool)


class CoveredSet:
    """Device-resident covered set (points + coverage voxels); Node::cloud_filtered / globalCloud."""

    def __init__(self, ctx, handle):
        self.ctx, self.handle = ctx, handle

    def free(self):
        if self.handle is not None:
            self.ctx._ck(self.ctx.L.asscpp_covered_free(self.ctx.h, self.handle))
            self.handle = None

    def clear(self):
        self.ctx._ck(self.ctx.L.asscpp_covered_clear(self.ctx.h, self.handle))

    def clone(self):
        out = C.c_int()
        self.ctx._ck(self.ctx.L.asscpp_covered_clone(self.ctx.h, self.handle, C.byref(out)))
        return CoveredSet(self.ctx, out.value)

    def or_set(self, other):
        self.ctx._ck(self.ctx.L.asscpp_covered_or_set(self.ctx.h, self.handle, other.handle))

    def or_views(self, poses, min_visible=-1):
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 7)
        self.ctx._ck(self.ctx.L.asscpp_covered_or_views(self.ctx.h, self.handle, _ptr(poses, C.c_double),
                                                        poses.shape[0], min_visible))

    def or_points(self, idx):
        idx = np.ascontiguousarray(idx, np.int32)
        self.ctx._ck(self.ctx.L.asscpp_covered_or_points(self.ctx.h, self.handle, _ptr(idx, C.c_int32), idx.shape[0]))

    def popcount(self):
        a, b = C.c_int64(), C.c_int64()
        self.ctx._ck(self.ctx.L.asscpp_covered_popcount(self.ctx.h, self.handle, C.byref(a), C.byref(b)))
        return a.value, b.value

    def export(self):
        pts = np.zeros((self.ctx.n_points + 63) // 64, np.uint64)
        vox = np.zeros((self.ctx.n_cov_voxels + 63) // 64, np.uint64)
        self.ctx._ck(self.ctx.L.asscpp_covered_export(self.ctx.h, self.handle, _ptr(pts, C.c_uint64),
                                                      _ptr(vox, C.c_uint64)))
        return pts, vox

    def load(self, point_bits, voxel_bits):
        pts = None if point_bits is None else np.ascontiguousarray(point_bits, np.uint64)
        vox = None if voxel_bits is None else np.ascontiguousarray(voxel_bits, np.uint64)
        self.ctx._ck(self.ctx.L.asscpp_covered_import(self.ctx.h, self.handle, _ptr(pts, C.c_uint64),
                                                      _ptr(vox, C.c_uint64)))

    def uncovered(self, pcl_aliasing=False):
        """pointCloudDiff (heuristic.cpp:88-145): (ids, centres) of the model's coverage voxels not in the set;
        pcl_aliasing=True reproduces the reference's unchecked linear-index lookup (quirk Q11)."""
        n_cov = self.ctx.n_cov_voxels
        ids = np.zeros(max(n_cov, 1), np.int32)
        ctr = np.zeros((max(n_cov, 1), 3), np.float32)
        n = C.c_int64()
        self.ctx._ck(self.ctx.L.asscpp_uncovered_voxels_ex(self.ctx.h, self.handle, 1 if pcl_aliasing else 0,
                                                           _ptr(ids, C.c_int32), _ptr(ctr, C.c_float), n_cov, C.byref(n)))
        return ids[:n.value].copy(), ctr[:n.value].copy()

    def uncovered_clusters(self, tolerance=0.8, min_size=5, max_size=50000):
        """clusteringPointCloud + findClusterBB (heuristic.cpp:147-246): (labels per coverage voxel, clusters)."""
        n_cov = self.ctx.n_cov_voxels
        labels = np.zeros(max(n_cov, 1), np.int32)
        out = np.zeros(max(n_cov, 1), CLUSTER_DTYPE)
        n = C.c_int64()
        self.ctx._ck(self.ctx.L.asscpp_uncovered_clusters(self.ctx.h, self.handle, tolerance, min_size, max_size,
                                                          _ptr(labels, C.c_int32), out.ctypes.data_as(C.c_void_p),
                                                          n_cov, C.byref(n)))
        return labels[:n_cov], out[:n.value].copy()

    def coverage_percent(self):
        out = C.c_float()
        self.ctx._ck(self.ctx.L.asscpp_coverage_percent(self.ctx.h, self.handle, C.byref(out)))
        return out.value


class CoverageContext:
    """One model cloud resident on one GPU (asscpp_create)."""

    def __init__(self, xyz, params=None, device=0):
        self.L = _abi.load()
        self.params = params or default_params()
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        self.h = C.c_void_p()
        rc = self.L.asscpp_create(device, _ptr(xyz, C.c_float), xyz.shape[0], C.byref(self.params), C.byref(self.h))
        if rc:
            self.h = C.c_void_p()
            raise AsscppError(rc, (self.L.asscpp_last_error(None) or b"").decode())
        info = ModelInfo()
        self._ck(self.L.asscpp_model_info(self.h, C.byref(info)))
        self.info = info
        self.n_points, self.n_occ_voxels, self.n_cov_voxels = info.n_points, info.n_occ_voxels, info.n_cov_voxels
        self.device = device

    def _ck(self, rc):
        if rc:
            raise AsscppError(rc, (self.L.asscpp_last_error(self.h) or b"").decode())

    def close(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self.L.asscpp_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- evaluation ------------------------------------------------------------------------------
    def eval_async(self, poses, covered=None, flags=0):
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 7)
        self._ck(self.L.asscpp_eval_batch(self.h, _ptr(poses, C.c_double), poses.shape[0],
                                          -1 if covered is None else covered.handle, flags))
        return poses.shape[0]

    def results(self, nviews):
        out = np.zeros(nviews, RESULT_DTYPE)
        self._ck(self.L.asscpp_get_results(self.h, out.ctypes.data_as(C.c_void_p), nviews))
        return out

    def eval_batch(self, poses, covered=None, flags=0):
        return self.results(self.eval_async(poses, covered, flags))

    def eval_groups(self, poses, group_size, covered=None, flags=0):
        """One A* expansion (astar.cpp:311-330): poses [n_candidates * group_size, 7], the sensors of a candidate
        consecutive.  gain_vox summed over a group = coverage voxels the candidate adds to `covered`."""
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 7)
        self._ck(self.L.asscpp_eval_groups(self.h, _ptr(poses, C.c_double), poses.shape[0], int(group_size),
                                           -1 if covered is None else covered.handle, flags))
        return self.results(poses.shape[0])

    def eval_ex(self, poses, covered=None, accum=None, accum_min_visible=5, group_size=1, flags=0, nviews=None):
        """asscpp_eval_ex: gains against `covered`, visible sets of views seeing more than accum_min_visible points
        OR-ed into `accum`, in one launch.  poses: host array [n, 7], or an int (device pointer) with nviews."""
        o = EvalOpts(-1 if covered is None else covered.handle, -1 if accum is None else accum.handle,
                     int(accum_min_visible), int(group_size), int(flags), 0, 0)
        if isinstance(poses, int):
            o.poses_on_device = 1
            self._ck(self.L.asscpp_eval_ex(self.h, C.c_void_p(poses), int(nviews), C.byref(o)))
            return int(nviews)
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 7)
        self._ck(self.L.asscpp_eval_ex(self.h, poses.ctypes.data_as(C.c_void_p), poses.shape[0], C.byref(o)))
        return poses.shape[0]

    def gains(self, nviews):
        """(gain_pts, gain_vox) of the last evaluation: 8 bytes per view device -> host."""
        gp, gv = np.zeros(nviews, np.uint32), np.zeros(nviews, np.uint32)
        self._ck(self.L.asscpp_get_gain(self.h, _ptr(gp, C.c_uint32), _ptr(gv, C.c_uint32), nviews))
        return gp, gv

    def accuracy(self, nviews):
        """Accuracy terms of the last evaluation issued with EVAL_ACCURACY (AsscppViewAccuracy per view)."""
        out = np.zeros(nviews, ACCURACY_DTYPE)
        self._ck(self.L.asscpp_get_accuracy(self.h, out.ctypes.data_as(C.c_void_p), nviews))
        return out

    def view_cells(self, nviews):
        """Per-view coverage-voxel sets of the last evaluation issued with EVAL_VIEW_CELLS: bool array
        [nviews, n_cov_voxels] (asscpp_get_view_cells)."""
        words = self.L.asscpp_view_cells_words(self.h)
        rows = np.zeros((nviews, max(words, 1)), np.uint32)
        self._ck(self.L.asscpp_get_view_cells(self.h, rows.ctypes.data_as(C.c_void_p), nviews))
        bits = np.unpackbits(rows.view(np.uint8), axis=1, bitorder="little")[:, :self.n_cov_voxels]
        return bits.astype(bool)

    def eval_device(self, d_poses_ptr, nviews, covered=None, flags=0):
        """poses already in HBM (raw device pointer); results stay on the device."""
        out = C.c_void_p()
        self._ck(self.L.asscpp_eval_batch_device(self.h, C.c_void_p(d_poses_ptr), nviews,
                                                 -1 if covered is None else covered.handle, flags, C.byref(out)))
        return out.value

    def get_visible(self, pose):
        pose = np.ascontiguousarray(pose, np.float64)
        idx = np.zeros(max(self.n_points, 1), np.int32)
        n = C.c_int64()
        self._ck(self.L.asscpp_get_visible(self.h, _ptr(pose, C.c_double), _ptr(idx, C.c_int32), idx.shape[0],
                                           C.byref(n)))
        return idx[: n.value].copy()

    def new_set(self):
        out = C.c_int()
        self._ck(self.L.asscpp_covered_new(self.h, C.byref(out)))
        return CoveredSet(self, out.value)

    def coverage_percent_cloud(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = C.c_float()
        self._ck(self.L.asscpp_coverage_percent_cloud(self.h, _ptr(xyz, C.c_float), xyz.shape[0], C.byref(out)))
        return out.value

    # -- streams / timing --------------------------------------------------------------------------
    def sync(self):
        self._ck(self.L.asscpp_sync(self.h))

    def stream(self):
        return self.L.asscpp_get_stream(self.h)

    def timer_begin(self):
        self._ck(self.L.asscpp_timer_begin(self.h))

    def timer_end(self):
        ms = C.c_float()
        self._ck(self.L.asscpp_timer_end(self.h, C.byref(ms)))
        return ms.value

    def eval_kernel_ms(self):
        ms, n = C.c_float(), C.c_int64()
        self._ck(self.L.asscpp_eval_kernel_ms(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def phase_cycles(self, enable=True):
        """Read + reset the per-phase SM-cycle counters (cull, grid, DDA, visit, views, CTA lifetime)."""
        out = np.zeros(16, np.uint64)
        self._ck(self.L.asscpp_debug_phase_cycles(self.h, int(enable), _ptr(out, C.c_uint64)))
        return out

    def launch_count(self):
        return self.L.asscpp_launch_count(self.h)

    # -- multi-GPU -----------------------------------------------------------------------------------
    @staticmethod
    def comm_unique_id():
        buf = np.zeros(128, np.uint8)
        rc = _abi.load().asscpp_comm_unique_id(_ptr(buf, C.c_uint8))
        if rc:
            raise AsscppError(rc, (_abi.load().asscpp_last_error(None) or b"").decode())
        return buf

    def comm_init(self, nranks, rank, uid):
        uid = np.ascontiguousarray(uid, np.uint8)
        self._ck(self.L.asscpp_comm_init(self.h, nranks, rank, _ptr(uid, C.c_uint8)))
        self.nranks, self.rank = nranks, rank

    def allgather_results_device(self, nviews_local):
        self._ck(self.L.asscpp_allgather_results(self.h, nviews_local, None))

    def allgather_results(self, nviews_local):
        out = np.zeros(nviews_local * self.nranks, RESULT_DTYPE)
        self._ck(self.L.asscpp_allgather_results(self.h, nviews_local, out.ctypes.data_as(C.c_void_p)))
        return out

    def allgather_gain(self, nviews_local, to_host=True):
        """gain-score allgather (8 bytes per view); to_host=False leaves the gathered scores on the device."""
        if not to_host:
            self._ck(self.L.asscpp_allgather_gain(self.h, nviews_local, None, None))
            return None
        gp = np.zeros(nviews_local * self.nranks, np.uint32)
        gv = np.zeros(nviews_local * self.nranks, np.uint32)
        self._ck(self.L.asscpp_allgather_gain(self.h, nviews_local, _ptr(gp, C.c_uint32), _ptr(gv, C.c_uint32)))
        return gp, gv

    def collective_ms(self):
        """device time of (allgather, OR-allreduce) between the last timer_begin / timer_end, with call counts."""
        a, b, na, nb = C.c_float(), C.c_float(), C.c_int64(), C.c_int64()
        self._ck(self.L.asscpp_collective_ms(self.h, C.byref(a), C.byref(na), C.byref(b), C.byref(nb)))
        return (a.value, na.value), (b.value, nb.value)

    def allreduce_or(self, covered):
        self._ck(self.L.asscpp_allreduce_or(self.h, covered.handle))

    def comm_destroy(self):
        self._ck(self.L.asscpp_comm_destroy(self.h))


class MultiContext:
    """One process, several GPUs (asscpp_multi_*): a context per device, batches sharded over them, covered sets
    mirrored and kept equal by the NCCL OR-allreduce."""

    def __init__(self, xyz, params=None, devices=(0,)):
        self.L = _abi.load()
        self.params = params or default_params()
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        dev = np.ascontiguousarray(devices, np.int32)
        self.h = C.c_void_p()
        rc = self.L.asscpp_multi_create(_ptr(dev, C.c_int32), len(dev), _ptr(xyz, C.c_float), xyz.shape[0],
                                        C.byref(self.params), C.byref(self.h))
        if rc:
            self.h = C.c_void_p()
            raise AsscppError(rc, (self.L.asscpp_last_error(None) or b"").decode())
        self.ndev = self.L.asscpp_multi_num_devices(self.h)

    def _ck(self, rc):
        if rc:
            raise AsscppError(rc, (self.L.asscpp_multi_last_error(self.h) or b"").decode())

    def close(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self.L.asscpp_multi_destroy(self.h)
            self.h = C.c_void_p()

    def new_set(self):
        out = C.c_int()
        self._ck(self.L.asscpp_multi_covered_new(self.h, C.byref(out)))
        return out.value

    def free_set(self, s):
        self._ck(self.L.asscpp_multi_covered_free(self.h, s))

    def popcount(self, s):
        a, b = C.c_int64(), C.c_int64()
        self._ck(self.L.asscpp_multi_covered_popcount(self.h, s, C.byref(a), C.byref(b)))
        return a.value, b.value

    def coverage_percent(self, s):
        out = C.c_float()
        self._ck(self.L.asscpp_multi_coverage_percent(self.h, s, C.byref(out)))
        return out.value

    def or_views(self, s, poses, min_visible=-1):
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 7)
        self._ck(self.L.asscpp_multi_covered_or_views(self.h, s, _ptr(poses, C.c_double), len(poses), min_visible))

    def eval(self, poses, covered=-1, accum=-1, accum_min_visible=5, group_size=1, flags=0):
        """(results, gain_pts, gain_vox) for all poses, evaluated on all devices."""
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 7)
        o = EvalOpts(int(covered), int(accum), int(accum_min_visible), int(group_size), int(flags), 0, 0)
        res = np.zeros(len(poses), RESULT_DTYPE)
        gp, gv = np.zeros(len(poses), np.uint32), np.zeros(len(poses), np.uint32)
        self._ck(self.L.asscpp_multi_eval(self.h, _ptr(poses, C.c_double), len(poses), C.byref(o),
                                          res.ctypes.data_as(C.c_void_p), _ptr(gp, C.c_uint32), _ptr(gv, C.c_uint32)))
        return res, gp, gv

    def adaptive_sampling(self, mesh, grid_start, grid_size, start_res=4.5, res_decrement=1.5, mounts=None,
                          global_set=None, cap=200000):
        """asscpp_adaptive_sampling (the batched C++ host loop of pathplanner.cpp:729-850):
        (robot poses [n, 7] newest first, sensor poses [n, sensors, 7], stages)."""
        from ._abi import STAGE_DTYPE, SamplingOpts
        mounts = mounts or (((0.0, 0.022, 0.065), (0.0, -0.349, 0.0)), ((0.0, 0.022, -0.065), (0.0, 0.349, 0.0)))
        o = SamplingOpts()
        o.grid_start[:] = list(grid_start)
        o.grid_size[:] = list(grid_size)
        o.start_res, o.res_decrement, o.orientation_res_deg = start_res, res_decrement, 45.0
        o.min_dist, o.max_dist, o.stop_uncovered_pct = 1.0, 4.0, 0.01
        o.n_sensors = len(mounts)
        for i, (off, rpy) in enumerate(mounts):
            # the reference keeps the mounts in float32 (Eigen::Vector3f, sspp/include/sspp/sensors.h:60)
            o.sensor_offset[i][:] = [float(np.float32(v)) for v in off]
            o.sensor_rpy[i][:] = [float(np.float32(v)) for v in rpy]
        gs = self.new_set() if global_set is None else global_set
        robot = np.zeros((cap, 7), np.float64)
        sens = np.zeros((cap, len(mounts), 7), np.float64)
        stages = np.zeros(16, STAGE_DTYPE)
        n, nst = C.c_int64(), C.c_int()
        self._ck(self.L.asscpp_adaptive_sampling(self.h, mesh.h, C.byref(o), gs, _ptr(robot, C.c_double),
                                                 _ptr(sens, C.c_double), cap, C.byref(n),
                                                 stages.ctypes.data_as(C.c_void_p), len(stages), C.byref(nst)))
        return robot[:n.value].copy(), sens[:n.value].copy(), stages[:nst.value].copy(), gs

    def export_set(self, s, device_index=0):
        """(point bits, voxel bits) of the set as device `device_index` holds it."""
        ctx = self.L.asscpp_multi_context(self.h, device_index)
        info = ModelInfo()
        self.L.asscpp_model_info(C.c_void_p(ctx), C.byref(info))
        pts = np.zeros((info.n_points + 63) // 64, np.uint64)
        vox = np.zeros((info.n_cov_voxels + 63) // 64, np.uint64)
        rc = self.L.asscpp_covered_export(C.c_void_p(ctx), s, _ptr(pts, C.c_uint64), _ptr(vox, C.c_uint64))
        if rc:
            raise AsscppError(rc, "covered_export")
        return pts, vox
