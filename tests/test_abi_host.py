"""CPU tests (-m "not gpu") of the C-ABI library: it loads, exports every symbol include/asscpp_gpu.h
declares, refuses to run without a GPU (no CPU fallback), and its host-only helpers (frustum planes,
robot->sensor transform, synthetic generators) agree with the oracle bit for bit."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def test_exports_every_declared_symbol(lib_built):
    from asscpp_b200 import _abi
    header = open(os.path.join(ROOT, "include", "asscpp_gpu.h")).read()
    declared = sorted(set(re.findall(r"\b(asscpp_[a-z0-9_]+)\s*\(", header)))
    assert declared == sorted(_abi.SYMBOLS)
    lib = C.CDLL(lib_built)
    for name in declared:
        assert hasattr(lib, name), name
    assert _abi.load().asscpp_abi_version() == 1


def test_result_struct_layout(lib_built):
    from asscpp_b200 import _abi
    assert _abi.RESULT_DTYPE.itemsize == 64
    assert _abi.RESULT_DTYPE.fields["cells"][1] == 32 and _abi.RESULT_DTYPE.fields["surv_hash"][1] == 56


def test_no_cpu_fallback(lib_built):
    """Without a CUDA device every compute entry fails loudly (ASSCPP_ENODEVICE), never silently on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import asscpp_b200 as A
    with pytest.raises(A.AsscppError) as e:
        A.CoverageContext(np.zeros((10, 3), np.float32))
    assert e.value.code == -2


def test_product_does_not_touch_oracle():
    """The oracle is test infrastructure: nothing under asscpp_b200/ or include/ may reference it."""
    bad = []
    for base in ("asscpp_b200", "include", "compat"):
        for dp, _, fs in os.walk(os.path.join(ROOT, base)):
            for f in fs:
                if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")) or f == "Makefile":
                    text = open(os.path.join(dp, f), errors="replace").read()
                    if re.search(r"oracle/|asscpp_oracle|from oracle|import oracle", text):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_planes_host_equals_oracle(lib_built, oracle_mod, space_2p5):
    import asscpp_b200 as A
    p = A.default_params()
    o = oracle_mod.Oracle(np.zeros((1, 3), np.float32))
    for pose in np.concatenate([space_2p5["cam0"][::31], space_2p5["cam1"][::29]]):
        pl, T = A.frustum_planes(p, pose)
        pl2, T2 = o.planes(pose)
        assert np.array_equal(pl.view(np.uint32), pl2.view(np.uint32))
        assert np.array_equal(T.view(np.uint32), T2.view(np.uint32))


def test_robot2sensor_equals_oracle_and_fixtures(lib_built, oracle_mod, space_2p5):
    import asscpp_b200 as A
    for k in range(0, 1225, 25):
        for cam, off, rpy in ((space_2p5["cam0"], (0, 0.022, 0.065), (0, -0.349, 0)),
                              (space_2p5["cam1"], (0, 0.022, -0.065), (0, 0.349, 0))):
            s = A.robot2sensor(space_2p5["uav"][k], off, rpy)
            assert np.array_equal(s, oracle_mod.robot2sensor(space_2p5["uav"][k], off, rpy))
            q = s[3:] if np.dot(s[3:], cam[k, 3:]) >= 0 else -s[3:]
            assert np.abs(s[:3] - cam[k, :3]).max() < 1e-4 and np.abs(q - cam[k, 3:]).max() < 6e-6


def test_synth_generators_deterministic(lib_built):
    import asscpp_b200 as A
    a = A.synth_structure(50_000)
    b = A.synth_structure(50_000)
    assert np.array_equal(a, b)
    assert a.min() >= 0.0 and a[:, 0].max() <= 40.0 and a[:, 1].max() <= 40.0 and a[:, 2].max() <= 10.0
    # a prefix of a longer stream is the same stream
    assert np.array_equal(A.synth_structure(1000), a[:1000])
    v = A.synth_views(2000)
    assert np.array_equal(v, A.synth_views(2000))
    assert (v[:, 2] >= 0.3).all()
    assert np.allclose(np.linalg.norm(v[:, 3:], axis=1), 1.0, atol=1e-12)
    # committed digest of the first 1000 points / 100 views (pins the generator across builds)
    assert int(np.frombuffer(a[:1000].tobytes(), np.uint32).sum(dtype=np.uint64)) == SYNTH_POINT_DIGEST
    assert int(np.frombuffer(np.round(v[:100] * 1e9).astype(np.int64).tobytes(), np.uint32).sum(dtype=np.uint64)) \
        == SYNTH_VIEW_DIGEST


SYNTH_POINT_DIGEST = 3175316270458
SYNTH_VIEW_DIGEST = 1637634120653


def test_viewport_transform_and_rpy_equal_oracle(lib_built, oracle_mod, etihad, space_2p5):
    """pointCloudViewportTransform / getRPY (occlusion_culling.cpp:538-618) on the host side of the C ABI are
    bit-identical to the oracle's restatement, and the transformed x is the depth along the camera's view axis."""
    import asscpp_b200 as A
    xyz, _ = etihad
    pts = xyz[::997]
    for pose in np.concatenate([space_2p5["cam0"][::97], space_2p5["cam1"][::89]]):
        assert np.array_equal(A.pose_rpy(pose), oracle_mod.pose_rpy(pose))
        a = A.viewport_transform(pose, pts)
        b = oracle_mod.viewport_transform(pose, pts)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        x, y, z, w = pose[3:]
        n = x * x + y * y + z * z + w * w
        view = np.array([1 - 2 * (y * y + z * z) / n, 2 * (x * y + w * z) / n, 2 * (x * z - w * y) / n])
        depth = (pts.astype(np.float64) - pose[:3]) @ view
        assert np.abs(a[:, 0] - depth).max() < 2e-5  # float32 stages over coordinates of tens of metres


def test_host_only_library_and_batch_transform(lib_built, space_2p5):
    """libasscpp_host.so carries the host helpers without CUDA; the batch robot->sensor transform equals the
    one-pose entry point bit for bit (sspp/src/sensors.cpp:47-90)."""
    import asscpp_b200 as A
    from asscpp_b200 import _abi
    assert os.path.exists(_abi.HOST_LIB_PATH)
    H = C.CDLL(_abi.HOST_LIB_PATH)
    for name in _abi.HOST_SYMBOLS:
        assert hasattr(H, name), name
    deps = os.popen("ldd %s" % _abi.HOST_LIB_PATH).read()
    assert "cuda" not in deps.lower() and "nccl" not in deps.lower(), deps
    uav = space_2p5["uav"][:200]
    off, rpy = (0.0, 0.022, 0.065), (0.0, -0.349, 0.0)
    batch = A.robot2sensor_batch(uav, off, rpy)
    single = np.stack([A.robot2sensor(u, off, rpy) for u in uav])
    assert np.array_equal(batch, single)


def test_reference_arm_never_maps_the_product_library(lib_built):
    """bench.py --impl reference: inputs from libasscpp_host.so, compute from oracle/; the CUDA library stays unmapped."""
    import json
    import subprocess
    import sys
    env = dict(os.environ, ASSCPP_BENCH_REPORT_MAPS="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "2",
                          "--points", "20000", "--views-per-step", "64", "--steps", "1", "--warmup", "3",
                          "--cpu-seconds", "1"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["value"] > 0
    maps = line["mapped_repo_libraries"]
    assert any("libasscpp_oracle" in m for m in maps) and any("libasscpp_host" in m for m in maps)
    assert not any("libasscpp_b200" in m for m in maps), maps
