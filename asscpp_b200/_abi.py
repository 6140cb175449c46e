"""ctypes view of include/asscpp_gpu.h (the C ABI of libasscpp_b200.so).

The library is built in-tree by asscpp_b200/csrc/Makefile (nvcc, sm_100a).  Loading fails loudly
when it is missing: there is no Python or CPU fallback for any compute entry point.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ASSCPP_LIB") or os.path.join(_HERE, "lib", "libasscpp_b200.so")

OK, EINVAL, ENODEVICE, ECUDA, ERANGE, ENOMEM, ESTATE, ENCCL = 0, -1, -2, -3, -4, -5, -6, -7
EVAL_HASH, EVAL_BRUTE, EVAL_FORCE_GLOBAL, EVAL_ACCURACY, EVAL_VIEW_CELLS = 1, 2, 4, 8, 16
VS_GRID_SHARED, VS_GRID_GLOBAL, VS_ORDERED = 1, 2, 4


class Params(C.Structure):
    _fields_ = [
        ("leaf", C.c_float),
        ("hfov_deg", C.c_float),
        ("vfov_deg", C.c_float),
        ("near_dist", C.c_float),
        ("far_dist", C.c_float),
        ("pcd_origin", C.c_float * 3),
        ("cov_leaf", C.c_float),
    ]


class ModelInfo(C.Structure):
    _fields_ = [
        ("n_points", C.c_int64),
        ("n_occ_voxels", C.c_int64),
        ("n_cov_voxels", C.c_int64),
        ("n_chunks", C.c_int64),
        ("min_b", C.c_int32 * 3),
        ("max_b", C.c_int32 * 3),
        ("sm_count", C.c_int32),
        ("reserved", C.c_int32),
    ]


class SamplingOpts(C.Structure):
    _fields_ = [
        ("grid_start", C.c_double * 3), ("grid_size", C.c_double * 3), ("start_res", C.c_double),
        ("res_decrement", C.c_double), ("orientation_res_deg", C.c_double), ("min_dist", C.c_double),
        ("max_dist", C.c_double), ("stop_uncovered_pct", C.c_double), ("n_sensors", C.c_int32), ("reserved", C.c_int32),
        ("sensor_offset", (C.c_double * 3) * 4), ("sensor_rpy", (C.c_double * 3) * 4),
    ]


STAGE_DTYPE = np.dtype([("res", np.float64), ("uncovered_pct_before", np.float64), ("uncovered_pct_after", np.float64),
                        ("grids", np.int64), ("candidates", np.int64), ("passed_mesh", np.int64), ("accepted", np.int64),
                        ("clusters", np.int64)])


class EvalOpts(C.Structure):
    _fields_ = [
        ("covered_set", C.c_int32),
        ("accum_set", C.c_int32),
        ("accum_min_visible", C.c_int64),
        ("group_size", C.c_int32),
        ("flags", C.c_uint32),
        ("poses_on_device", C.c_uint32),
        ("reserved", C.c_uint32),
    ]


# numpy mirror of AsscppViewResult
RESULT_DTYPE = np.dtype(
    [
        ("survivors", np.uint32),
        ("visible", np.uint32),
        ("carry", np.uint32),
        ("unique_rays", np.uint32),
        ("gain_pts", np.uint32),
        ("gain_vox", np.uint32),
        ("status", np.uint32),
        ("reserved", np.uint32),
        ("cells", np.uint64),
        ("steps", np.uint64),
        ("vis_hash", np.uint64),
        ("surv_hash", np.uint64),
    ]
)
assert RESULT_DTYPE.itemsize == 64

# numpy mirror of AsscppViewAccuracy
ACCURACY_DTYPE = np.dtype(
    [
        ("err_sum_depth", np.float64),
        ("err_sum_z", np.float64),
        ("depth_min", np.float32),
        ("depth_max", np.float32),
        ("visible", np.uint32),
        ("reserved", np.uint32),
    ]
)
assert ACCURACY_DTYPE.itemsize == 32

# numpy mirror of AsscppCluster
CLUSTER_DTYPE = np.dtype(
    [
        ("size", np.int32),
        ("first_id", np.int32),
        ("min_ijk", np.int32, 3),
        ("max_ijk", np.int32, 3),
        ("grid_start", np.float64, 3),
        ("grid_size", np.float64, 3),
    ]
)
assert CLUSTER_DTYPE.itemsize == 80

# every symbol include/asscpp_gpu.h declares (checked by tests/test_abi_symbols.py)
SYMBOLS = [
    "asscpp_create", "asscpp_destroy", "asscpp_last_error", "asscpp_abi_version", "asscpp_model_info",
    "asscpp_default_params", "asscpp_eval_batch", "asscpp_get_results", "asscpp_get_counts", "asscpp_get_gain",
    "asscpp_eval_batch_device", "asscpp_get_visible", "asscpp_covered_new", "asscpp_covered_free",
    "asscpp_covered_clear", "asscpp_covered_clone", "asscpp_covered_or_set", "asscpp_covered_or_views",
    "asscpp_covered_or_points", "asscpp_covered_popcount", "asscpp_covered_export", "asscpp_covered_import",
    "asscpp_coverage_percent", "asscpp_coverage_percent_cloud", "asscpp_sync", "asscpp_get_stream",
    "asscpp_timer_begin", "asscpp_timer_end", "asscpp_launch_count", "asscpp_eval_kernel_ms",
    "asscpp_debug_phase_cycles",
    "asscpp_comm_unique_id", "asscpp_comm_init", "asscpp_allgather_results", "asscpp_allreduce_or",
    "asscpp_comm_destroy", "asscpp_synth_structure_v1", "asscpp_synth_views_v1", "asscpp_robot2sensor",
    "asscpp_frustum_planes", "asscpp_get_accuracy", "asscpp_pose_rpy", "asscpp_viewport_transform",
    "asscpp_uncovered_voxels", "asscpp_uncovered_voxels_ex", "asscpp_uncovered_clusters", "asscpp_uncovered_ids_noalloc", "asscpp_view_cells_words", "asscpp_get_view_cells",
    "asscpp_eval_groups", "asscpp_mesh_create", "asscpp_mesh_destroy", "asscpp_mesh_num_triangles", "asscpp_mesh_launch_count",
    "asscpp_mesh_last_error", "asscpp_mesh_filter", "asscpp_mesh_segments_free",
    "asscpp_covered_or_cloud", "asscpp_covered_invert_voxels", "asscpp_comm_init_all",
    "asscpp_multi_create", "asscpp_multi_destroy", "asscpp_multi_num_devices", "asscpp_multi_context",
    "asscpp_multi_last_error", "asscpp_multi_covered_new", "asscpp_multi_covered_free", "asscpp_multi_covered_clear",
    "asscpp_multi_covered_popcount", "asscpp_multi_coverage_percent", "asscpp_multi_allreduce_or", "asscpp_multi_eval",
    "asscpp_multi_covered_or_views", "asscpp_grid_candidates", "asscpp_adaptive_sampling",
    "asscpp_eval_ex", "asscpp_allgather_gain", "asscpp_collective_ms", "asscpp_robot2sensor_batch",
]

# the host-only subset, also exported by lib/libasscpp_host.so (no CUDA inside)
HOST_LIB_PATH = os.path.join(_HERE, "lib", "libasscpp_host.so")
HOST_SYMBOLS = ["asscpp_synth_structure_v1", "asscpp_synth_views_v1", "asscpp_robot2sensor", "asscpp_robot2sensor_batch"]

_lib = None


def load():
    """dlopen the in-tree library and declare the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libasscpp_b200.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'` or "
            "`make -C asscpp_b200/csrc`. There is no CPU fallback." % LIB_PATH
        )
    L = C.CDLL(LIB_PATH)
    vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
    f32p, f64p, i32p, u32p, u64p, u8p, i64p = (
        C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_uint32),
        C.POINTER(C.c_uint64), C.POINTER(C.c_uint8), C.POINTER(C.c_int64))
    sig = {
        "asscpp_create": (i32, [i32, f32p, i64, C.POINTER(Params), C.POINTER(vp)]),
        "asscpp_destroy": (i32, [vp]),
        "asscpp_last_error": (C.c_char_p, [vp]),
        "asscpp_abi_version": (i32, []),
        "asscpp_model_info": (i32, [vp, C.POINTER(ModelInfo)]),
        "asscpp_default_params": (None, [C.POINTER(Params)]),
        "asscpp_eval_batch": (i32, [vp, f64p, i64, i32, C.c_uint32]),
        "asscpp_eval_groups": (i32, [vp, f64p, i64, i32, i32, C.c_uint32]),
        "asscpp_get_results": (i32, [vp, vp, i64]),
        "asscpp_get_counts": (i32, [vp, u32p, u32p, u32p, i64]),
        "asscpp_get_gain": (i32, [vp, u32p, u32p, i64]),
        "asscpp_eval_batch_device": (i32, [vp, vp, i64, i32, C.c_uint32, C.POINTER(vp)]),
        "asscpp_get_visible": (i32, [vp, f64p, i32p, i64, i64p]),
        "asscpp_covered_new": (i32, [vp, C.POINTER(i32)]),
        "asscpp_covered_free": (i32, [vp, i32]),
        "asscpp_covered_clear": (i32, [vp, i32]),
        "asscpp_covered_clone": (i32, [vp, i32, C.POINTER(i32)]),
        "asscpp_covered_or_set": (i32, [vp, i32, i32]),
        "asscpp_covered_or_views": (i32, [vp, i32, f64p, i64, i64]),
        "asscpp_covered_or_points": (i32, [vp, i32, i32p, i64]),
        "asscpp_covered_popcount": (i32, [vp, i32, i64p, i64p]),
        "asscpp_covered_export": (i32, [vp, i32, u64p, u64p]),
        "asscpp_covered_import": (i32, [vp, i32, u64p, u64p]),
        "asscpp_coverage_percent": (i32, [vp, i32, f32p]),
        "asscpp_coverage_percent_cloud": (i32, [vp, f32p, i64, f32p]),
        "asscpp_sync": (i32, [vp]),
        "asscpp_get_stream": (vp, [vp]),
        "asscpp_timer_begin": (i32, [vp]),
        "asscpp_timer_end": (i32, [vp, f32p]),
        "asscpp_launch_count": (i64, [vp]),
        "asscpp_eval_kernel_ms": (i32, [vp, f32p, i64p]),
        "asscpp_debug_phase_cycles": (i32, [vp, i32, u64p]),
        "asscpp_comm_unique_id": (i32, [u8p]),
        "asscpp_comm_init": (i32, [vp, i32, i32, u8p]),
        "asscpp_allgather_results": (i32, [vp, i64, vp]),
        "asscpp_allreduce_or": (i32, [vp, i32]),
        "asscpp_comm_destroy": (i32, [vp]),
        "asscpp_synth_structure_v1": (i32, [C.c_uint64, C.c_uint64, i64, f32p]),
        "asscpp_synth_views_v1": (i32, [C.c_uint64, C.c_uint64, i64, f64p]),
        "asscpp_robot2sensor": (i32, [f64p, f64p, f64p, f64p]),
        "asscpp_frustum_planes": (i32, [C.POINTER(Params), f64p, f32p, f32p]),
        "asscpp_get_accuracy": (i32, [vp, vp, i64]),
        "asscpp_view_cells_words": (i64, [vp]),
        "asscpp_get_view_cells": (i32, [vp, vp, i64]),
        "asscpp_pose_rpy": (i32, [f64p, f64p]),
        "asscpp_viewport_transform": (i32, [f64p, f32p, i64, f32p]),
        "asscpp_mesh_create": (i32, [i32, f32p, i64, i32p, i64, C.POINTER(vp)]),
        "asscpp_mesh_destroy": (i32, [vp]),
        "asscpp_mesh_num_triangles": (i64, [vp]),
        "asscpp_mesh_launch_count": (i64, [vp]),
        "asscpp_mesh_last_error": (C.c_char_p, [vp]),
        "asscpp_mesh_filter": (i32, [vp, f64p, i64, f64p, C.c_double, C.c_double, i32p, f64p, u8p]),
        "asscpp_mesh_segments_free": (i32, [vp, f64p, f64p, i64, u8p]),
        "asscpp_uncovered_voxels": (i32, [vp, i32, i32p, f32p, i64, i64p]),
        "asscpp_uncovered_voxels_ex": (i32, [vp, i32, C.c_uint32, i32p, f32p, i64, i64p]),
        "asscpp_uncovered_ids_noalloc": (i32, [vp, i64, vp, i64, vp, vp, C.c_char_p, i64]),
        "asscpp_uncovered_clusters": (i32, [vp, i32, C.c_double, C.c_int32, C.c_int32, i32p, vp, i64, i64p]),
        "asscpp_comm_init_all": (i32, [C.POINTER(vp), i32]),
        "asscpp_multi_create": (i32, [i32p, i32, f32p, i64, C.POINTER(Params), C.POINTER(vp)]),
        "asscpp_multi_destroy": (i32, [vp]),
        "asscpp_multi_num_devices": (i32, [vp]),
        "asscpp_multi_context": (vp, [vp, i32]),
        "asscpp_multi_last_error": (C.c_char_p, [vp]),
        "asscpp_multi_covered_new": (i32, [vp, C.POINTER(i32)]),
        "asscpp_multi_covered_free": (i32, [vp, i32]),
        "asscpp_multi_covered_clear": (i32, [vp, i32]),
        "asscpp_multi_covered_popcount": (i32, [vp, i32, i64p, i64p]),
        "asscpp_multi_coverage_percent": (i32, [vp, i32, f32p]),
        "asscpp_multi_allreduce_or": (i32, [vp, i32]),
        "asscpp_multi_eval": (i32, [vp, f64p, i64, C.POINTER(EvalOpts), vp, u32p, u32p]),
        "asscpp_multi_covered_or_views": (i32, [vp, i32, f64p, i64, i64]),
        "asscpp_grid_candidates": (i64, [f64p, f64p, C.c_double, C.c_double, f64p, i64]),
        "asscpp_adaptive_sampling": (i32, [vp, vp, C.POINTER(SamplingOpts), i32, f64p, f64p, i64, i64p, vp, i32,
                                           C.POINTER(i32)]),
        "asscpp_covered_or_cloud": (i32, [vp, i32, f32p, i64]),
        "asscpp_covered_invert_voxels": (i32, [vp, i32]),
        "asscpp_eval_ex": (i32, [vp, vp, i64, C.POINTER(EvalOpts)]),
        "asscpp_allgather_gain": (i32, [vp, i64, u32p, u32p]),
        "asscpp_collective_ms": (i32, [vp, f32p, i64p, f32p, i64p]),
        "asscpp_robot2sensor_batch": (i32, [f64p, i64, f64p, f64p, f64p]),
    }
    assert sorted(sig) == sorted(SYMBOLS)
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L
