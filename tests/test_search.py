"""SSPP search around the hot path (BASELINE.json config 1): search-space formats, node
connection, A* with batched child scoring.  CPU: host logic with the oracle scorer.  GPU: the same search driven
through asscpp_eval_groups equals the oracle-driven one step for step."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT

sys.path.insert(0, os.path.join(ROOT, "tests"))
from asscpp_b200 import search as SR  # noqa: E402
from backends import OracleScorer  # noqa: E402

START = (3.0, -34.5, 9.0)  # coverage_heuristic_test.cpp:83


def small_space(space_2p5, rows):
    return SR.SearchSpace.from_poses(space_2p5["uav"][rows], np.stack([space_2p5["cam0"][rows], space_2p5["cam1"][rows]], 1))


def test_insert_is_head_insertion_without_duplicates(space_2p5):
    rows = [5, 6, 7, 6, 5, 8]
    s = small_space(space_2p5, rows)
    assert s.n == 4
    assert np.array_equal(s.robots, space_2p5["uav"][[8, 7, 6, 5]])   # list order = reverse insertion order
    assert s.ids.tolist() == [3, 2, 1, 0]
    assert np.array_equal(s.sensors[0, 1], space_2p5["cam1"][8])


def test_connect_matches_double_loop(space_2p5):
    rows = np.arange(300, 420)
    s = small_space(space_2p5, rows)
    free = lambda a, b: (np.floor(a[:, 0] + b[:, 0]).astype(int) % 3) != 0  # any symmetric stand-in for the mesh
    total = s.connect(2.5, free, block=50)
    R = s.robots
    want_total = 0
    for t in range(s.n):                      # connectNodes, pathplanner.cpp:534-568
        want = []
        for S in range(s.n):
            d = SR.dist(R[S], R[t])
            if d <= 2.5 and S != t:
                if (R[S, :3] != R[t, :3]).any():
                    if free(R[t:t + 1, :3], R[S:S + 1, :3])[0]:
                        want.append(S)
                else:
                    want.append(S)
        assert s.children[t].tolist() == want
        want_total += len(want)
    assert total == want_total == s.n_connections > 0
    # same position, other yaw: always connected
    same = [(t, c) for t in range(s.n) for c in s.children[t] if (R[t, :3] == R[c, :3]).all()]
    assert same


def test_wire_formats_round_trip(tmp_path, space_2p5, oracle_mod):
    rows = np.arange(600, 660)
    s = small_space(space_2p5, rows)
    s.connect(2.5)
    f = [str(tmp_path / n) for n in ("uav.txt", "cam0.txt", "cam1.txt")]
    SR.save_regular_grid(s, *f)
    t = SR.load_regular_grid(*f)
    assert t.n == s.n
    np.testing.assert_allclose(t.robots, s.robots, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(t.sensors, s.sensors, rtol=1e-5, atol=1e-6)
    # connection file: 14 numbers per edge, 3 decimals (pathplanner.cpp:269-299)
    path = str(tmp_path / "space.txt")
    s.save(path)
    lines = open(path).read().splitlines()
    assert len(lines) == s.n_connections and all(len(l.split()) == 14 for l in lines)
    assert all(len(x.split(".")[1]) == 3 for x in lines[0].split())
    from asscpp_b200.adaptive import MOUNTS
    u = SR.SearchSpace.load(path, oracle_mod.robot2sensor, MOUNTS)
    assert u.n_connections == s.n_connections
    # every edge survives with poses rounded to 3 decimals
    Rs, Ru = s.robots, u.robots
    es = sorted((tuple(np.round(Rs[i], 3)), tuple(np.round(Rs[c], 3))) for i in range(s.n) for c in s.children[i])
    eu = sorted((tuple(np.round(Ru[i], 3)), tuple(np.round(Ru[c], 3))) for i in range(u.n) for c in u.children[i])
    assert np.allclose(np.array(es), np.array(eu), atol=1.1e-3)
    # sensors recomputed from the robot pose (sensors.cpp:47-90)
    i = 3
    assert np.array_equal(u.sensors[i, 0], oracle_mod.robot2sensor(Ru[i], *MOUNTS[0]))


def test_llist_order():
    class N:
        def __init__(self, f, tag):
            self.f, self.tag = f, tag
    lst = []
    for f, tag in ((1.0, "a"), (3.0, "b"), (2.0, "c"), (3.0, "d"), (0.5, "e")):
        SR._llist_add(lst, N(f, tag))
    assert [n.tag for n in lst] == ["d", "b", "c", "a", "e"]  # newer node ahead of equal f (llist.cpp:96-109)


@pytest.mark.parametrize("accumulate", [False, True])
def test_astar_with_oracle_scorer(oracle_mod, etihad, space_2p5, accumulate):
    xyz, _ = etihad
    rows = np.arange(560, 760)
    s = small_space(space_2p5, rows)
    s.connect(2.5)
    o = oracle_mod.Oracle(xyz)
    sc = OracleScorer(o, oracle_mod, accumulate)
    start = s.robots[40]
    root = min(i for i in range(s.n) if (s.robots[i, :3] == start[:3]).all())  # findRoot: first nearest in list order
    target = sc.score(None, s.sensors[root][None])[0][0] + (0.8 if not accumulate else 5.0)
    out = SR.astar_search(s, start, sc, target, 0.5, max_expansions=2000)
    assert out is not None and out["path"][0] == root and len(out["path"]) > 1
    # the path is connected, distances add up, the last node meets the target
    R = s.robots
    dsum = 0.0
    for a, b in zip(out["path"][:-1], out["path"][1:]):
        assert b in s.children[a]
        dsum += SR.dist(R[a], R[b])
    assert out["distance"] == pytest.approx(dsum, rel=1e-12)
    assert target - out["coverage"][-1] <= 0.5
    if accumulate:
        assert all(b >= a for a, b in zip(out["coverage"][:-1], out["coverage"][1:]))  # covered set only grows
    else:
        # quirk Q10: a node's coverage is its own view's coverage
        last = out["path"][-1]
        own = sc.score(None, s.sensors[last][None])[0][0]
        assert out["coverage"][-1] == own


@pytest.mark.gpu
@pytest.mark.parametrize("accumulate", [False, True])
def test_astar_gpu_equals_oracle(oracle_mod, etihad, space_2p5, accumulate):
    """Config 1: the coverage search over the bundled model and the 1 225-waypoint space, children scored by ONE
    asscpp_eval_groups launch per expansion, against the literal per-child oracle scoring: identical paths,
    coverages and f values (floats compared exactly)."""
    import asscpp_b200 as A
    xyz, _ = etihad
    g = np.load(os.path.join(ROOT, "tests", "golden", "etihad_mesh.npz"))
    mesh = A.Mesh(g["vertices"], g["triangles"])
    s = SR.SearchSpace.from_poses(space_2p5["uav"], np.stack([space_2p5["cam0"], space_2p5["cam1"]], 1))
    assert s.n == 1225
    s.connect(2.5, mesh.segments_free)   # coverage_heuristic_test.cpp:87 connection radius
    t = SR.SearchSpace.from_poses(space_2p5["uav"], np.stack([space_2p5["cam0"], space_2p5["cam1"]], 1))
    t.connect(2.5, oracle_mod.Mesh(g["vertices"], g["triangles"]).segments_free)
    assert s.n_connections == t.n_connections and all(np.array_equal(a, b) for a, b in zip(s.children, t.children))
    ctx = A.CoverageContext(xyz, A.default_params())
    target, tol = (5.0, 1.0) if not accumulate else (20.0, 1.0)  # :108 target 5 +- 1
    gs = SR.GpuScorer(ctx, accumulate)
    got = SR.astar_search(s, START, gs, target, tol, max_expansions=3000)
    want = SR.astar_search(t, START, OracleScorer(oracle_mod.Oracle(xyz), oracle_mod, accumulate), target, tol,
                           max_expansions=3000)
    assert got is not None and want is not None
    assert got["path"] == want["path"] and got["coverage"] == want["coverage"] and got["f"] == want["f"]
    assert got["expanded"] == want["expanded"] and got["created"] == want["created"]
    assert gs.launches == got["expanded"]   # root + one launch per expansion that had children (last one terminates)
    ctx.close(); mesh.close()


@pytest.mark.gpu
@pytest.mark.parametrize("model", ["etihad", "synthetic_fine"])
@pytest.mark.parametrize("split", [True, False])
def test_eval_groups_union_gain(oracle_mod, etihad, space_2p5, model, split, monkeypatch):
    """asscpp_eval_groups: the gain_vox sum of a group = |voxels(union of the group's visible clouds) \\ covered|;
    every other field equals asscpp_eval_batch's.  Both schedules: a group per CTA (large batches) and split-group
    mode (batches of up to 4096 views: every view its own work unit, group_gain_kernel afterwards); per view the two
    give the same numbers."""
    import asscpp_b200 as A
    O = oracle_mod
    monkeypatch.setenv("ASSCPP_SPLIT_GROUP_MAX", "4096" if split else "0")
    if model == "etihad":
        xyz = etihad[0]
        prm, oprm = A.default_params(), None
        rows = np.arange(600, 720)
        poses = np.stack([space_2p5["cam0"][rows], space_2p5["cam1"][rows], space_2p5["cam0"][rows + 1]], 1).reshape(-1, 7)
        seed_poses = space_2p5["cam0"][560:600]
    else:
        xyz = A.synth_structure(300000)
        prm, oprm = A.default_params(leaf=0.05, cov_leaf=0.5), O.make_params(leaf=0.05, cov_leaf=0.5)
        v = A.synth_views(300)
        poses, seed_poses = v[:240], v[240:]
    ctx = A.CoverageContext(xyz, prm)
    o = O.Oracle(xyz, oprm)
    covered = ctx.new_set()
    covered.or_views(seed_poses, -1)
    cv = np.zeros(o.vox_words(), np.uint64)
    cp = np.zeros(o.pts_words(), np.uint64)
    o.accumulate(seed_poses, cp, cv, min_visible=-1)
    in_cov = np.unpackbits(cv.view(np.uint8), bitorder="little").astype(bool)
    ids = o.cov_ids()
    for group in (3, 2, 1):
        n = (len(poses) // group) * group
        res = ctx.eval_groups(poses[:n], group, covered=covered)
        plain = ctx.eval_batch(poses[:n], covered=covered)
        for f in ("survivors", "visible", "carry", "unique_rays", "gain_pts", "cells", "steps"):
            assert np.array_equal(res[f], plain[f]), f
        got = res["gain_vox"].astype(np.int64).reshape(-1, group).sum(1)
        vox = [np.unique(ids[o.eval_view(p)["visible"]]) for p in poses[:n]]
        for gi in range(n // group):
            u = np.unique(np.concatenate(vox[gi * group:(gi + 1) * group]))
            assert got[gi] == int((~in_cov[u]).sum()), (group, gi)
        if group == 1:
            assert np.array_equal(res["gain_vox"], plain["gain_vox"])
        if split:
            monkeypatch.setenv("ASSCPP_SPLIT_GROUP_MAX", "0")
            other = A.CoverageContext(xyz, prm)
            monkeypatch.setenv("ASSCPP_SPLIT_GROUP_MAX", "4096")
            oc = other.new_set()
            oc.load(*covered.export())
            assert np.array_equal(other.eval_groups(poses[:n], group, covered=oc)["gain_vox"], res["gain_vox"]), group
            other.close()
    assert got.sum() > 0
    # ragged tail (the last group is short), a group larger than the batch, empty batch, bad group size
    r7 = ctx.eval_groups(poses[:7], 3, covered=covered)
    assert np.array_equal(r7["gain_vox"][:6], ctx.eval_groups(poses[:6], 3, covered=covered)["gain_vox"])
    assert r7["gain_vox"][6] == ctx.eval_batch(poses[6:7], covered=covered)["gain_vox"][0]
    big = ctx.eval_groups(poses[:5], 64, covered=covered)
    u = np.unique(np.concatenate([np.unique(ids[o.eval_view(p)["visible"]]) for p in poses[:5]]))
    assert int(big["gain_vox"].sum()) == int((~in_cov[u]).sum())
    assert len(ctx.eval_groups(poses[:0], 2)) == 0
    with pytest.raises(A.AsscppError):
        ctx.eval_groups(poses[:4], 0)
    ctx.close()


@pytest.mark.gpu
def test_small_batch_single_launch_equals_general_path(lib_built, etihad, space_2p5, monkeypatch):
    """Small batches (the A* use, at most 512 views) are ONE launch: poses as kernel parameters, planes computed by
    the CTA that takes the view, group gains by the last view of the group, gains mirrored into mapped host memory
    (asscpp_b200/csrc/abi.cu run_eval / finalize_small).  Every field equals the general multi-launch path's, with
    and without groups, accuracy terms included; a batch holding a quirk-Q5 view falls back to the general path."""
    import asscpp_b200 as A
    xyz = etihad[0]
    monkeypatch.setenv("ASSCPP_NO_SMALL", "1")
    general = A.CoverageContext(xyz, A.default_params())
    monkeypatch.delenv("ASSCPP_NO_SMALL")
    small = A.CoverageContext(xyz, A.default_params())
    cov_g, cov_s = general.new_set(), small.new_set()
    seed = space_2p5["cam0"][600:640]
    cov_g.or_views(seed, 5); cov_s.or_views(seed, 5)
    fields = ("survivors", "visible", "carry", "unique_rays", "cells", "steps", "gain_pts", "gain_vox", "vis_hash",
              "surv_hash")
    for n, group in ((1, 1), (2, 2), (7, 1), (40, 2), (63, 3), (64, 1), (66, 2), (400, 2), (512, 8), (512, 1)):
        k = (n + 1) // 2
        poses = np.stack([space_2p5["cam0"][500:500 + k], space_2p5["cam1"][500:500 + k]], 1).reshape(-1, 7)[:n]
        for flags in (A.EVAL_HASH, A.EVAL_HASH | A.EVAL_ACCURACY):
            out = []
            for ctx, cov in ((general, cov_g), (small, cov_s)):
                before = ctx.launch_count()
                ctx.eval_ex(poses, covered=cov, group_size=group, flags=flags)
                gp, gv = ctx.gains(n)
                res = ctx.results(n)
                acc = ctx.accuracy(n) if flags & A.EVAL_ACCURACY else None
                out.append((gp, gv, res, acc, ctx.launch_count() - before))
            (gp0, gv0, r0, a0, l0), (gp1, gv1, r1, a1, l1) = out
            assert l1 == 1 and l0 >= 3, (n, group, l0, l1)
            assert np.array_equal(gp0, gp1) and np.array_equal(gv0, gv1), (n, group)
            assert np.array_equal(gp1, r1["gain_pts"]) and np.array_equal(gv1, r1["gain_vox"])
            for f in fields:
                assert np.array_equal(r0[f], r1[f]), (n, group, f)
            if a0 is not None:
                assert a0.tobytes() == a1.tobytes(), (n, group)
        assert r1["visible"].sum() > 0 or n < 7
    general.close(); small.close()
    # a quirk-Q5 view (tests/test_gpu_parity.py::test_carry_inside_large_batch) in a small batch
    rng = np.random.default_rng(5)
    xyz = np.concatenate([np.array([(42, 6, 2), (14, 2, 2), (2.5, 2, 2), (14, 3, 6)], np.float32),
                          (rng.random((5000, 3)) * np.array([60, 30, 12]) + np.array([0, -40, 0])).astype(np.float32)])
    pa = A.default_params(leaf=4.0, far_dist=60.0, cov_leaf=2.0)
    poses = np.array([[1, 2, 2, 0, 0, 0, 1.0], [1, -20, 2, 0, 0, 0, 1.0], [5, -30, 6, 0, 0, 0.38268343, 0.92387953],
                      [1, 2, 2, 0, 0, 0, 1.0]])
    monkeypatch.setenv("ASSCPP_NO_SMALL", "1")
    general = A.CoverageContext(xyz, pa)
    monkeypatch.delenv("ASSCPP_NO_SMALL")
    small = A.CoverageContext(xyz, pa)
    for group in (1, 2):
        general.eval_ex(poses, group_size=group, flags=A.EVAL_HASH)
        small.eval_ex(poses, group_size=group, flags=A.EVAL_HASH)
        g0, g1 = general.gains(4), small.gains(4)
        r0, r1 = general.results(4), small.results(4)
        assert np.array_equal(g0[0], g1[0]) and np.array_equal(g0[1], g1[1])
        for f in fields + ("status",):
            assert np.array_equal(r0[f], r1[f]), (group, f)
        assert r1["carry"][0] > 0 and (r1["status"][0] & A.VS_ORDERED)
    general.close(); small.close()
