"""GPU parity tests (-m gpu): the sm_100a path, called through the C ABI, against the CPU oracle.

Bar: bit-exact.  Every integer the path produces -- survivor / visible counts, carry events, unique
rays, grid cells, DDA step counts, point and voxel gains, and 64-bit hashes of the survivor and visible
index sets -- must equal the oracle's on the same inputs; visible index LISTS are compared verbatim on
sampled views.  Epsilon classes E1-E5 are counted by the oracle and reported; no
mismatch is tolerated inside them either.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, assert_results_equal

pytestmark = pytest.mark.gpu

NTHREADS = os.cpu_count() or 1


@pytest.fixture(scope="module")
def A(lib_built):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import asscpp_b200
    return asscpp_b200


@pytest.fixture(scope="module")
def etihad_ctx(A, etihad):
    ctx = A.CoverageContext(etihad[0], A.default_params())
    yield ctx
    ctx.close()


@pytest.fixture(scope="module")
def etihad_oracle(oracle_mod, etihad):
    return oracle_mod.Oracle(etihad[0])


def test_bundled_model_all_fixture_views(A, oracle_mod, etihad_ctx, etihad_oracle, space_2p5):
    """Config 1: bundled model x the 2 x 1 225 sensor poses of cscpp/txt, reference defaults."""
    poses = np.concatenate([space_2p5["cam0"], space_2p5["cam1"]])
    res = etihad_ctx.eval_batch(poses, flags=A.EVAL_HASH)
    st = etihad_oracle.eval_batch(poses, memo=True, nthreads=NTHREADS)
    assert_results_equal(oracle_mod, res, st, "etihad 2.5m")
    assert int(res["visible"].sum()) > 500000
    assert (res["status"] & 0x80000000 == 0).all()


def test_bundled_model_large_search_space(A, oracle_mod, etihad_ctx, etihad_oracle, space_1p5):
    """Config 1: the 2 x 17 583 poses of sspp/resources (1.5 m search space)."""
    poses = np.concatenate([space_1p5["cam0"], space_1p5["cam1"]])
    res = etihad_ctx.eval_batch(poses, flags=A.EVAL_HASH)
    st = etihad_oracle.eval_batch(poses, memo=True, nthreads=NTHREADS)
    assert_results_equal(oracle_mod, res, st, "etihad 1.5m")


def test_golden_fixture(A, oracle_mod, etihad_ctx, space_2p5):
    """Committed golden vectors (tools/make_golden.py), no oracle run involved."""
    g = np.load(os.path.join(GOLDEN, "etihad_oracle_golden.npz"))
    poses = np.concatenate([space_2p5["cam0"], space_2p5["cam1"]])[g["view_index"]]
    res = etihad_ctx.eval_batch(poses, flags=A.EVAL_HASH)
    assert_results_equal(oracle_mod, res, g["stats"], "golden")
    vis = etihad_ctx.get_visible(poses[int(g["list_view"])])
    assert np.array_equal(vis, g["visible_list"])


def test_visible_lists_verbatim(A, etihad_ctx, etihad_oracle, space_2p5):
    """extractVisibleSurface returns points in ascending original index order (occlusion_culling.cpp:317,371)."""
    poses = np.concatenate([space_2p5["cam0"][600:700:7], space_2p5["cam1"][300:900:41]])
    for pose in poses:
        want = etihad_oracle.eval_view(pose)["visible"]
        got = etihad_ctx.get_visible(pose)
        assert np.array_equal(got, want)


@pytest.mark.parametrize("flags_name", ["brute", "global_grid"])
def test_alternate_paths_agree(A, oracle_mod, etihad_ctx, etihad_oracle, space_2p5, flags_name):
    """Bucket pre-cull off (every point streamed per view) and the bit-grid kept in global memory
    instead of shared memory: same bits."""
    fl = A.EVAL_HASH | (A.EVAL_BRUTE if flags_name == "brute" else A.EVAL_FORCE_GLOBAL)
    poses = np.concatenate([space_2p5["cam0"][::3], space_2p5["cam1"][1::3]])
    res = etihad_ctx.eval_batch(poses, flags=fl)
    st = etihad_oracle.eval_batch(poses, memo=True, nthreads=NTHREADS)
    assert_results_equal(oracle_mod, res, st, flags_name)
    want = A.VS_GRID_GLOBAL if flags_name == "global_grid" else A.VS_GRID_SHARED
    assert ((res["status"] & 3 == want) | (res["survivors"] == 0)).all()


@pytest.mark.parametrize("leaf,cov_leaf", [(0.5, 0.5), (0.05, 0.5), (0.05, 0.05), (0.2, 0.3)])
def test_synthetic_structure(A, oracle_mod, leaf, cov_leaf):
    """synth_structure_v1 at a size the oracle finishes in seconds; 0.05 m is the grid of configs 3/5."""
    xyz = A.synth_structure(400_000)
    poses = A.synth_views(600)
    ctx = A.CoverageContext(xyz, A.default_params(leaf=leaf, cov_leaf=cov_leaf))
    orc = oracle_mod.Oracle(xyz, oracle_mod.make_params(leaf=leaf, cov_leaf=cov_leaf))
    assert ctx.n_cov_voxels == orc.n_cov and ctx.n_occ_voxels == orc.n_occ
    res = ctx.eval_batch(poses, flags=A.EVAL_HASH)
    st = orc.eval_batch(poses, memo=True, nthreads=NTHREADS)
    assert_results_equal(oracle_mod, res, st, "synth leaf %g cov %g" % (leaf, cov_leaf))
    for v in np.nonzero(st[:, oracle_mod.ST_VISIBLE] > 0)[0][:3]:
        assert np.array_equal(ctx.get_visible(poses[v]), orc.eval_view(poses[v])["visible"])
    ctx.close()


def test_config2_full_size_bit_exact(A, oracle_mod):
    """Config 2: 10k candidate viewpoints vs a synthetic 1M-point structure, single B200, bit-exact
    visible-set check (hash of every visible index set + counts) against the oracle."""
    xyz = A.synth_structure(1_000_000)
    poses = A.synth_views(10_000)
    ctx = A.CoverageContext(xyz, A.default_params())
    orc = oracle_mod.Oracle(xyz, oracle_mod.make_params())
    res = ctx.eval_batch(poses, flags=A.EVAL_HASH)
    st = orc.eval_batch(poses, memo=True, nthreads=NTHREADS)
    assert_results_equal(oracle_mod, res, st, "config 2")
    ctx.close()


def test_epsilon_classes_reported(A, oracle_mod, etihad_ctx, etihad_oracle, space_2p5):
    """North star: points within epsilon of a frustum plane / voxel boundary are counted and reported.
    They are compared like every other point: the visible lists still match verbatim."""
    total = np.zeros(5, np.int64)
    for pose in space_2p5["cam0"][600:660:6]:
        r = etihad_oracle.eval_view(pose, want_eps=True)
        total += r["eps"]
        assert np.array_equal(etihad_ctx.get_visible(pose), r["visible"])
    print("epsilon classes E1..E5 over 10 views:", total.tolist())
    assert total[0] > 0  # some points do sit within 1e-4 of a plane


def test_zero_direction_components(A, oracle_mod):
    """Quirk Q8: rays with exactly zero direction components (t_delta = inf, t_max = +-inf) and the x -> y -> z
    `<=` cascade on infinities (voxel_grid_occlusion_estimation.cpp:513-519, 565-579).  A lattice of points at
    exact cell centres and cameras at exact cell centres / faces, looking along each axis: whole planes of targets
    share a coordinate with the origin.  The device step loop evaluates the cascade without its last comparison
    (kernels.cu: dda_burst), which this scene would expose if the stated invariant did not hold."""
    for leaf, near in ((0.5, 0.7), (0.25, 0.3), (1.0, 0.1)):
        g = (np.arange(-8, 9, dtype=np.float32) + 0.5) * np.float32(leaf)
        X, Y, Z = np.meshgrid(g, g, g, indexing="ij")
        keep = ((np.abs(X) + np.abs(Y) + np.abs(Z)) / leaf).astype(np.int64) % 3 != 0   # holes: free and occluded rays
        xyz = np.stack([X[keep], Y[keep], Z[keep]], 1).astype(np.float32)
        pa = A.default_params(leaf=leaf, cov_leaf=leaf, near_dist=near, far_dist=near + 12 * leaf)
        po = oracle_mod.make_params(leaf=leaf, cov_leaf=leaf, near=near, far=near + 12 * leaf)
        c = float(np.float32(0.5 * leaf))
        s = float(np.sqrt(0.5))
        quats = [(0, 0, 0, 1.0), (0, 0, 1.0, 0), (0, 0, s, s), (0, 0, -s, s), (0, s, 0, s), (0, -s, 0, s)]  # +-x +-y +-z
        poses = []
        for q in quats:
            for o in ((c, c, c), (0.0, c, c), (c, 0.0, c), (c, c, 0.0), (-c, c, -c), (c + 4 * leaf, c, c - 2 * leaf)):
                poses.append(o + q)
        poses = np.array(poses, np.float64)
        ctx = A.CoverageContext(xyz, pa)
        orc = oracle_mod.Oracle(xyz, po)
        for flags in (A.EVAL_HASH, A.EVAL_HASH | A.EVAL_FORCE_GLOBAL):
            res = ctx.eval_batch(poses, flags=flags)
            st = orc.eval_batch(poses, memo=True)   # step counts with the early exit at the first occluder
            assert_results_equal(oracle_mod, res, st, "lattice leaf %g flags %d" % (leaf, flags))
            lit = orc.eval_batch(poses, memo=False)  # the literal per-point loop: same sets and counts
            for col in (oracle_mod.ST_VISIBLE, oracle_mod.ST_SURVIVORS):
                assert np.array_equal(lit[:, col], st[:, col])
        assert res["visible"].sum() > 0 and (res["visible"] < res["survivors"]).any()
        ctx.close()


def test_sticky_state_carry_path(A, oracle_mod):
    """Quirk Q5 (occlusion_culling.cpp:308,332-335): same scene as tests/test_oracle.py; the GPU hands
    such views to ordered_views_kernel, which replays the reference loop in original index order."""
    pa = A.default_params(leaf=4.0, far_dist=60.0, cov_leaf=4.0)
    po = oracle_mod.make_params(leaf=4.0, far=60.0, cov_leaf=4.0)
    pose = np.array([1, 2, 2, 0, 0, 0, 1.0])
    pts = {"A": (42, 6, 2), "B": (14, 2, 2), "C": (2.5, 2, 2), "E": (14, 3, 6)}
    for order in ("ABCE", "BCAE", "EABC", "CEBA"):
        xyz = np.array([pts[ch] for ch in order], np.float32)
        ctx = A.CoverageContext(xyz, pa)
        orc = oracle_mod.Oracle(xyz, po)
        res = ctx.eval_batch(np.stack([pose, pose]), flags=A.EVAL_HASH)
        st = orc.eval_batch(np.stack([pose, pose]), memo=False)
        cols = [(n, c) for n, c in [("survivors", 0), ("visible", 1), ("carry", 2), ("unique_rays", 3), ("cells", 4),
                                    ("gain_pts", 6), ("gain_vox", 7), ("vis_hash", 8), ("surv_hash", 9)]]
        for name, col in cols:
            assert np.array_equal(res[name].astype(np.uint64), st[:, col].astype(np.uint64)), (order, name)
        assert (res["status"] & A.VS_ORDERED != 0).all() and (res["carry"] == 2).all()
        assert np.array_equal(ctx.get_visible(pose), orc.eval_view(pose)["visible"])
        ctx.close()


def test_carry_inside_large_batch(A, oracle_mod, etihad):
    """A carry view mixed into a normal batch: only that view takes the ordered path."""
    rng = np.random.default_rng(5)
    xyz = np.concatenate([np.array([(42, 6, 2), (14, 2, 2), (2.5, 2, 2), (14, 3, 6)], np.float32),
                          (rng.random((5000, 3)) * np.array([60, 30, 12]) + np.array([0, -40, 0])).astype(np.float32)])
    pa = A.default_params(leaf=4.0, far_dist=60.0, cov_leaf=2.0)
    po = oracle_mod.make_params(leaf=4.0, far=60.0, cov_leaf=2.0)
    poses = np.array([[1, 2, 2, 0, 0, 0, 1.0], [1, -20, 2, 0, 0, 0, 1.0], [5, -30, 6, 0, 0, 0.38268343, 0.92387953],
                      [1, 2, 2, 0, 0, 0, 1.0]])
    ctx = A.CoverageContext(xyz, pa)
    orc = oracle_mod.Oracle(xyz, po)
    res = ctx.eval_batch(poses, flags=A.EVAL_HASH)
    st = orc.eval_batch(poses, memo=False)
    for name, col in [("survivors", 0), ("visible", 1), ("carry", 2), ("gain_pts", 6), ("gain_vox", 7), ("vis_hash", 8)]:
        assert np.array_equal(res[name].astype(np.uint64), st[:, col].astype(np.uint64)), name
    assert res["carry"][0] > 0 and res["carry"][3] > 0
    ctx.close()


def test_covered_sets_and_gain(A, oracle_mod, etihad_ctx, etihad_oracle, space_2p5):
    """Gain = popcount of newly visible points / coverage voxels against a running covered set
    (coverage_path_planning_heuristic.cpp:473-483 in bitset form), union with the >5 rule (:344-346)."""
    o = etihad_oracle
    poses = np.concatenate([space_2p5["cam0"][560:700], space_2p5["cam1"][560:700]])
    cp = np.zeros(o.pts_words(), np.uint64)
    cv = np.zeros(o.vox_words(), np.uint64)
    o.accumulate(poses[:100], cp, cv, min_visible=5, nthreads=NTHREADS)
    s = etihad_ctx.new_set()
    s.or_views(poses[:100], min_visible=5)
    gp, gv = s.export()
    assert np.array_equal(gp, cp) and np.array_equal(gv, cv)
    assert s.popcount() == (int(np.unpackbits(cp.view(np.uint8)).sum()), int(np.unpackbits(cv.view(np.uint8)).sum()))
    assert np.float32(s.coverage_percent()) == np.float32(o.coverage_percent_vox(cv))
    res = etihad_ctx.eval_batch(poses, covered=s, flags=A.EVAL_HASH)
    st = o.eval_batch(poses, memo=True, nthreads=NTHREADS, covered_pts=cp, covered_vox=cv)
    assert_results_equal(oracle_mod, res, st, "gain vs covered set")
    # import/export round trip, clone, or_set, or_points
    t = etihad_ctx.new_set()
    t.load(cp, cv)
    assert all(np.array_equal(a, b) for a, b in zip(t.export(), (cp, cv)))
    u = s.clone()
    u.or_set(t)
    assert all(np.array_equal(a, b) for a, b in zip(u.export(), (cp, cv)))
    w = etihad_ctx.new_set()
    vis = o.eval_view(poses[120])["visible"]
    w.or_points(vis)
    wp, wv = w.export()
    assert int(np.unpackbits(wp.view(np.uint8)).sum()) == len(vis)
    assert np.float32(w.coverage_percent()) == np.float32(o.coverage_percent_idx(vis))
    for x in (s, t, u, w):
        x.free()


def test_eval_ex_scores_and_accumulates_in_one_launch(A, oracle_mod, etihad_ctx, etihad_oracle, space_2p5):
    """asscpp_eval_ex: gains against one set while the visible sets of views seeing more than 5 points are OR-ed into
    another (heuristic.cpp:342-346 + :473-483 in one pass); asscpp_get_gain moves the same two numbers as the
    64-byte records hold."""
    o = etihad_oracle
    poses = np.concatenate([space_2p5["cam0"][560:660], space_2p5["cam1"][560:660]])
    cp = np.zeros(o.pts_words(), np.uint64)
    cv = np.zeros(o.vox_words(), np.uint64)
    o.accumulate(poses[:60], cp, cv, min_visible=5, nthreads=NTHREADS)
    cov = etihad_ctx.new_set()
    cov.load(cp, cv)
    acc = etihad_ctx.new_set()
    n = etihad_ctx.eval_ex(poses, covered=cov, accum=acc, accum_min_visible=5, flags=A.EVAL_HASH)
    gp, gv = etihad_ctx.gains(n)
    res = etihad_ctx.results(n)
    st = o.eval_batch(poses, memo=True, nthreads=NTHREADS, covered_pts=cp, covered_vox=cv)
    assert_results_equal(oracle_mod, res, st, "eval_ex")
    assert np.array_equal(gp, res["gain_pts"]) and np.array_equal(gv, res["gain_vox"])
    ap = np.zeros(o.pts_words(), np.uint64)
    av = np.zeros(o.vox_words(), np.uint64)
    o.accumulate(poses, ap, av, min_visible=5, nthreads=NTHREADS)
    gp2, gv2 = acc.export()
    assert np.array_equal(gp2, ap) and np.array_equal(gv2, av)
    # the accumulator may not be the set scored against
    with pytest.raises(A.AsscppError):
        etihad_ctx.eval_ex(poses[:4], covered=cov, accum=cov)
    # grouped form through the same entry point
    n = etihad_ctx.eval_ex(poses, covered=cov, group_size=2)
    grp = etihad_ctx.results(n)
    ref = etihad_ctx.eval_groups(poses, 2, covered=cov)
    assert np.array_equal(grp["gain_vox"], ref["gain_vox"])
    cov.free()
    acc.free()


def test_coverage_percent_of_arbitrary_cloud(A, etihad_ctx, etihad_oracle, etihad, space_2p5):
    """calcCoveragePercent(cloud) voxelises any cloud, e.g. VoxelGrid centroids (heuristic.cpp:476-482)."""
    xyz = etihad[0]
    vis = etihad_oracle.eval_view(space_2p5["cam0"][610])["visible"]
    pct = etihad_ctx.coverage_percent_cloud(xyz[vis])
    assert np.float32(pct) == np.float32(etihad_oracle.coverage_percent_idx(vis))
    # points outside the model grid match nothing
    assert etihad_ctx.coverage_percent_cloud(np.array([[1e3, 1e3, 1e3]], np.float32)) == 0.0
    assert etihad_ctx.coverage_percent_cloud(np.zeros((0, 3), np.float32)) == 0.0
    assert np.float32(etihad_ctx.coverage_percent_cloud(xyz)) == np.float32(100.0)


def test_edge_cases(A, oracle_mod):
    """Empty model, single point, zero views, blind views, degenerate quaternion."""
    ctx = A.CoverageContext(np.zeros((0, 3), np.float32), A.default_params())
    res = ctx.eval_batch(np.array([[0, 0, 0, 0, 0, 0, 1.0]]))
    assert res["survivors"][0] == 0 and res["visible"][0] == 0
    assert len(ctx.eval_batch(np.zeros((0, 7)))) == 0
    ctx.close()
    one = np.array([[3.1, 0.2, 0.1]], np.float32)
    ctx = A.CoverageContext(one, A.default_params())
    orc = oracle_mod.Oracle(one)
    poses = np.array([[0, 0, 0, 0, 0, 0, 1.0], [0, 0, 0, 0, 0, 1.0, 0.0], [0, 0, 0, 0, 0, 0, 0.0],
                      [100, 100, 100, 0, 0, 0, 1.0]])
    res = ctx.eval_batch(poses, flags=A.EVAL_HASH)
    st = orc.eval_batch(poses, memo=True)
    assert_results_equal(oracle_mod, res, st, "single point")
    assert list(res["visible"]) == [1, 0, 0, 0]
    ctx.close()
    with pytest.raises(A.AsscppError):
        A.CoverageContext(np.array([[np.nan, 0, 0]], np.float32), A.default_params())
    with pytest.raises(A.AsscppError):  # <PCL> "Leaf size is too small"
        A.CoverageContext(np.array([[0, 0, 0], [4000, 4000, 4000]], np.float32), A.default_params(leaf=0.001))


def test_idempotent_and_order_independent(A, etihad_ctx, space_2p5):
    """Size-independent properties: re-evaluation is bit-identical; a batch is a set of independent views."""
    poses = np.concatenate([space_2p5["cam0"][400:900], space_2p5["cam1"][400:900]])
    a = etihad_ctx.eval_batch(poses, flags=A.EVAL_HASH)
    b = etihad_ctx.eval_batch(poses, flags=A.EVAL_HASH)
    assert a.tobytes() == b.tobytes()
    perm = np.random.default_rng(1).permutation(len(poses))
    c = etihad_ctx.eval_batch(poses[perm], flags=A.EVAL_HASH)
    assert c.tobytes() == a[perm].tobytes()


def test_full_size_properties_10m(A):
    """BASELINE config 3 size on one GPU (10M points, 0.05 m): properties that need no oracle --
    bucketed cull == brute-force cull, shared-memory grid == global grid, union popcount consistency."""
    xyz = A.synth_structure(10_000_000)
    poses = A.synth_views(100_000)[37::400]
    ctx = A.CoverageContext(xyz, A.default_params(leaf=0.05))
    a = ctx.eval_batch(poses, flags=A.EVAL_HASH)
    b = ctx.eval_batch(poses, flags=A.EVAL_HASH | A.EVAL_BRUTE)
    c = ctx.eval_batch(poses, flags=A.EVAL_HASH | A.EVAL_FORCE_GLOBAL)
    drop = ["status"]
    names = [n for n in a.dtype.names if n not in drop]
    for n in names:
        assert np.array_equal(a[n], b[n]), n
        assert np.array_equal(a[n], c[n]), n
    assert a["visible"].sum() > 0 and (a["visible"] <= a["survivors"]).all() and (a["gain_pts"] == a["visible"]).all()
    s = ctx.new_set()
    s.or_views(poses, min_visible=-1)
    npts, nvox = s.popcount()
    assert 0 < npts <= int(a["visible"].sum()) and npts >= int(a["visible"].max())
    again = ctx.eval_batch(poses, covered=s)
    assert (again["gain_pts"] == 0).all() and (again["gain_vox"] == 0).all()
    ctx.close()


def test_config3_parity_subset_10m(A, oracle_mod):
    """BASELINE config 3 (10M points, 100k-view pool, 0.05 m grid): a random 1 024-view subset against the oracle,
    every integer field and both set hashes bit-exact."""
    xyz = A.synth_structure(10_000_000)
    pool = A.synth_views(100_000)
    sub = pool[np.random.default_rng(3).choice(len(pool), 1024, replace=False)]
    ctx = A.CoverageContext(xyz, A.default_params(leaf=0.05, cov_leaf=0.5))
    orc = oracle_mod.Oracle(xyz, oracle_mod.make_params(leaf=0.05, cov_leaf=0.5))
    res = ctx.eval_batch(sub, flags=A.EVAL_HASH)
    st = orc.eval_batch(sub, memo=True, nthreads=NTHREADS)
    assert_results_equal(oracle_mod, res, st, "config 3 subset")
    assert res["visible"].sum() > 0
    ctx.close()


def test_config5_parity_subset_50m(A, oracle_mod):
    """BASELINE config 5 size (50M points, 0.05 m grid): 256 views of the sweep against the oracle."""
    xyz = A.synth_structure(50_000_000)
    sub = A.synth_views(1_000_000)[np.random.default_rng(5).choice(1_000_000, 256, replace=False)]
    ctx = A.CoverageContext(xyz, A.default_params(leaf=0.05, cov_leaf=0.5))
    orc = oracle_mod.Oracle(xyz, oracle_mod.make_params(leaf=0.05, cov_leaf=0.5))
    res = ctx.eval_batch(sub, flags=A.EVAL_HASH)
    st = orc.eval_batch(sub, memo=True, nthreads=NTHREADS)
    assert_results_equal(oracle_mod, res, st, "config 5 subset")
    assert res["visible"].sum() > 0
    ctx.close()


@pytest.mark.parametrize("with_mesh", [True, False])
def test_config4_adaptive_loop(A, oracle_mod, etihad, with_mesh):
    """Config 4: the ASSCPP adaptive loop (pathplanner.cpp:729-850 skeleton) driven by the CUDA path and by
    the oracle: identical candidate lists, identical covered bitset and acceptance count per iteration.
    with_mesh: samples filtered by the structure's triangle mesh (heuristic.cpp:317-335, asscpp_mesh_filter);
    without: the voxel stand-off fallback used for models that have no mesh."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from asscpp_b200 import adaptive as AD
    from backends import OracleBackend
    xyz, _ = etihad
    model = AD.VoxelModel(xyz, 0.5)
    ctx = A.CoverageContext(xyz, A.default_params())
    assert model.n == ctx.n_cov_voxels
    gmesh = omesh = None
    if with_mesh:
        g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "etihad_mesh.npz"))
        gmesh, omesh = A.Mesh(g["vertices"], g["triangles"]), oracle_mod.Mesh(g["vertices"], g["triangles"])
    tg = AD.adaptive_loop(AD.GpuBackend(ctx, mesh=gmesh), model, (0, 0, 4.5), (36, 50, 9))
    to = AD.adaptive_loop(OracleBackend(oracle_mod.Oracle(xyz), oracle_mod, NTHREADS, mesh=omesh), model, (0, 0, 4.5),
                          (36, 50, 9))
    assert len(tg) == len(to) >= 3
    for a, b in zip(tg, to):
        assert a["candidates"] == b["candidates"] and a["accepted"] == b["accepted"]
        assert np.array_equal(a["bits"], b["bits"])
        assert a["uncovered_pct"] == b["uncovered_pct"] and a["clusters"] == b["clusters"]
    assert tg[-1]["uncovered_pct"] < 1.0
    if with_mesh:
        assert [t["candidates"] for t in tg[:3]] == [480, 184, 472]
        gmesh.close()
    ctx.close()


@pytest.mark.gpu
def test_nccl_legs_two_gpus():
    """asscpp_allgather_results / asscpp_allreduce_or over real NCCL (needs >= 2 GPUs; skipped on a 1-GPU box)."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29617",
                          os.path.join(root, "tools", "nccl_check.py")], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "OK" in out.stdout, out.stdout + out.stderr


@pytest.mark.gpu
def test_accuracy_terms(A, oracle_mod, etihad_ctx, etihad_oracle, etihad, space_2p5):
    """§8f rank 1: calcAvgAccuracy (both overloads) and the depth extrema SSMaxMinAccuracy folds
    (occlusion_culling.cpp:445-520) per view.  Depths are float32 and bit-exact (host libm feeds the same
    sines / cosines to both sides); the GPU sums are the exact sums of the per-point double terms rounded once,
    the reference's left-to-right double sum may differ from that by at most visible/2 ulp: tolerance
    visible * 2^-53 relative, written here."""
    O = oracle_mod
    xyz, _ = etihad
    poses = np.ascontiguousarray(np.concatenate([space_2p5["cam0"][400:900:3], space_2p5["cam1"][400:900:3]]))
    res = etihad_ctx.eval_batch(poses, flags=A.EVAL_ACCURACY | A.EVAL_HASH)
    acc = etihad_ctx.accuracy(len(poses))
    assert np.array_equal(acc["visible"], res["visible"])
    checked = 0
    for v, pose in enumerate(poses):
        r = etihad_oracle.eval_view(pose)
        vis = r["visible"]
        assert len(vis) == int(res["visible"][v])
        if len(vis) == 0:
            assert acc["err_sum_depth"][v] == 0 and acc["err_sum_z"][v] == 0
            assert np.isposinf(acc["depth_min"][v]) and np.isneginf(acc["depth_max"][v])
            continue
        sd, sz, lo, hi = etihad_oracle.accuracy_idx(pose, vis)
        tol = len(vis) * 2.0 ** -53
        assert abs(acc["err_sum_depth"][v] - sd) <= tol * sd
        assert abs(acc["err_sum_z"][v] - sz) <= tol * sz
        assert np.float32(lo) == acc["depth_min"][v] and np.float32(hi) == acc["depth_max"][v]
        # the GPU value is the exact sum: compare against exact rational arithmetic on the oracle's depths
        if checked < 6:
            from fractions import Fraction
            t = O.viewport_transform(pose, xyz[vis])
            ex = sum(Fraction(0.0000285 * float(d) * float(d)) for d in t[:, 0])
            assert acc["err_sum_depth"][v] == float(ex)
        checked += 1
    assert checked > 50
    # run-to-run determinism of the integer accumulation
    etihad_ctx.eval_batch(poses, flags=A.EVAL_ACCURACY)
    acc2 = etihad_ctx.accuracy(len(poses))
    assert acc.tobytes() == acc2.tobytes()
    # SSMaxMinAccuracy over this search space (occlusion_culling.cpp:486-520), including its initial values
    mx, mn = np.finfo(np.float64).tiny, np.finfo(np.float64).max
    for v in range(len(poses)):
        if acc["visible"][v]:
            mx = max(mx, float(acc["depth_max"][v]))
            mn = min(mn, float(acc["depth_min"][v]))
    assert 0.7 < mn < mx < 7.5


@pytest.mark.gpu
def test_accuracy_terms_synthetic_fine_grid(A, oracle_mod):
    """Same bar on the synthetic structure at 0.05 m (inside-chunk voxels, cut voxels, shared and global grids)."""
    O = oracle_mod
    xyz = A.synth_structure(300_000)
    poses = A.synth_views(96)
    ctx = A.CoverageContext(xyz, A.default_params(leaf=0.05, cov_leaf=0.5))
    orc = O.Oracle(xyz, O.make_params(leaf=0.05, cov_leaf=0.5))
    for flags in (A.EVAL_ACCURACY, A.EVAL_ACCURACY | A.EVAL_FORCE_GLOBAL, A.EVAL_ACCURACY | A.EVAL_BRUTE):
        res = ctx.eval_batch(poses, flags=flags)
        acc = ctx.accuracy(len(poses))
        for v, pose in enumerate(poses):
            vis = orc.eval_view(pose)["visible"]
            assert len(vis) == int(res["visible"][v]) == int(acc["visible"][v])
            if len(vis) == 0:
                continue
            sd, sz, lo, hi = orc.accuracy_idx(pose, vis)
            tol = len(vis) * 2.0 ** -53
            assert abs(acc["err_sum_depth"][v] - sd) <= tol * sd and abs(acc["err_sum_z"][v] - sz) <= tol * sz + 1e-300
            assert np.float32(lo) == acc["depth_min"][v] and np.float32(hi) == acc["depth_max"][v]
    ctx.close()


@pytest.mark.gpu
def test_uncovered_regions(A, oracle_mod, etihad_ctx, etihad_oracle, space_2p5):
    """§8f rank 2: pointCloudDiff, clusteringPointCloud and findClusterBB (heuristic.cpp:88-246) on the device
    against the oracle's literal restatement (itself pinned against scipy in tests/test_oracle.py)."""
    O = oracle_mod
    poses = np.ascontiguousarray(np.concatenate([space_2p5["cam0"][::7], space_2p5["cam1"][::7]]))
    s = etihad_ctx.new_set()
    for cut in (0, 40, len(poses)):  # nothing covered, a little covered, the whole sample
        s.clear()
        cv = np.zeros(etihad_oracle.vox_words(), np.uint64)
        cp = np.zeros(etihad_oracle.pts_words(), np.uint64)
        if cut:
            s.or_views(poses[:cut], 5)
            etihad_oracle.accumulate(poses[:cut], cp, cv, min_visible=5, nthreads=NTHREADS)
        ids, ctr = s.uncovered()
        ids_o, ctr_o = etihad_oracle.uncovered(cv)
        assert np.array_equal(ids, ids_o)
        assert np.array_equal(ctr.view(np.uint32), ctr_o.view(np.uint32))
        labels, clusters = s.uncovered_clusters(0.8, 5, 50000)
        lab_o, k, ties = O.euclidean_clusters(ctr_o, 0.8, 5, 50000)
        assert ties == 0 and len(clusters) == k
        # same partition and same sizes, largest first; ties in size may be ordered differently
        mine = labels[ids]
        assert np.array_equal(mine >= 0, lab_o >= 0)
        pairs = set(zip(mine[mine >= 0].tolist(), lab_o[lab_o >= 0].tolist()))
        assert len(pairs) == k == len(set(a for a, _ in pairs)) == len(set(b for _, b in pairs))
        assert np.all(np.diff(clusters["size"]) <= 0)
        assert np.array_equal(clusters["size"], np.bincount(lab_o[lab_o >= 0], minlength=k)[:k]) or k == 0
        assert np.all(labels[np.setdiff1d(np.arange(etihad_ctx.n_cov_voxels), ids)] == -1)
        for c in range(k):
            start, size = O.cluster_bb(ctr_o[mine == c], 0.5)
            assert np.array_equal(clusters["grid_start"][c], start) and np.array_equal(clusters["grid_size"][c], size)
            assert clusters["first_id"][c] == ids[mine == c].min()
    s.free()


@pytest.mark.gpu
def test_uncovered_with_reference_index_aliasing(A, oracle_mod, etihad_ctx, etihad_oracle, space_2p5):
    """asscpp_uncovered_voxels_ex(ASSCPP_DIFF_PCL_ALIASING): pointCloudDiff's unchecked linear-index lookup (quirk Q11,
    heuristic.cpp:120-123) on the device equals the oracle's restatement (which equals the reference's own
    pointCloudDiff, tests/test_ref_pin.py)."""
    o = etihad_oracle
    for lo, hi in ((560, 600), (600, 700), (300, 900)):
        poses = np.concatenate([space_2p5["cam0"][lo:hi], space_2p5["cam1"][lo:hi]])
        cp = np.zeros(o.pts_words(), np.uint64)
        cv = np.zeros(o.vox_words(), np.uint64)
        o.accumulate(poses, cp, cv, min_visible=5, nthreads=NTHREADS)
        s = etihad_ctx.new_set()
        s.load(cp, cv)
        for alias in (False, True):
            ids, ctr = s.uncovered(pcl_aliasing=alias)
            ids_o, ctr_o = o.uncovered(cv, pcl_aliasing=alias)
            assert np.array_equal(ids, ids_o) and np.array_equal(ctr, ctr_o), (lo, hi, alias)
        s.free()
    e = etihad_ctx.new_set()  # empty set: nothing to alias into
    assert len(e.uncovered(pcl_aliasing=True)[0]) == o.n_cov
    e.free()


@pytest.mark.gpu
def test_single_process_multi_gpu_evaluator(A, oracle_mod, etihad, space_2p5):
    """asscpp_multi_*: one process, every visible GPU (1 on the test box, 2-8 under gpurun --gpus N): a batch sharded
    over the devices gives the results of one context, and the mirrored covered set equals the oracle's union on
    every device."""
    import torch
    ndev = torch.cuda.device_count()
    o = oracle_mod.Oracle(etihad[0])
    poses = np.concatenate([space_2p5["cam0"][500:701], space_2p5["cam1"][500:701]])  # 402 views: uneven shards
    m = A.MultiContext(etihad[0], A.default_params(), devices=list(range(ndev)))
    cov, acc = m.new_set(), m.new_set()
    m.or_views(cov, poses[:60], 5)
    cp = np.zeros(o.pts_words(), np.uint64)
    cv = np.zeros(o.vox_words(), np.uint64)
    o.accumulate(poses[:60], cp, cv, min_visible=5, nthreads=NTHREADS)
    for d in range(ndev):
        gp, gv = m.export_set(cov, d)
        assert np.array_equal(gp, cp) and np.array_equal(gv, cv), d
    res, gpts, gvox = m.eval(poses, covered=cov, accum=acc, accum_min_visible=5, flags=A.EVAL_HASH)
    st = o.eval_batch(poses, memo=True, nthreads=NTHREADS, covered_pts=cp, covered_vox=cv)
    assert_results_equal(oracle_mod, res, st, "multi eval")
    assert np.array_equal(gpts, res["gain_pts"]) and np.array_equal(gvox, res["gain_vox"])
    ap = np.zeros(o.pts_words(), np.uint64)
    av = np.zeros(o.vox_words(), np.uint64)
    o.accumulate(poses, ap, av, min_visible=5, nthreads=NTHREADS)
    for d in range(ndev):
        gp, gv = m.export_set(acc, d)
        assert np.array_equal(gp, ap) and np.array_equal(gv, av), d
    # grouped (the sensors of a candidate stay on one device)
    res2, _, _ = m.eval(poses, covered=cov, group_size=2)
    one = A.CoverageContext(etihad[0], A.default_params())
    s1 = one.new_set()
    s1.load(cp, cv)
    ref = one.eval_groups(poses, 2, covered=s1)
    assert np.array_equal(res2["gain_vox"], ref["gain_vox"])
    assert np.float32(m.coverage_percent(cov)) == np.float32(o.coverage_percent_vox(cv))
    one.close()
    m.close()


@pytest.mark.gpu
def test_adaptive_sampling_host_loop(A, etihad):
    """asscpp_adaptive_sampling: the C++ host loop of PathPlanner::dynamicNodesGenerationAndConnection
    (sspp/src/pathplanner.cpp:729-850), batched -- one mesh launch and one evaluation launch per stage over every
    visible GPU -- produces, row for row and in the search space's order, the waypoints the reference's OWN loop
    produced with the reference's own CPU heuristic (tests/golden/dsscpp_reference_run.npz, tools/reproduce_dsscpp.py)."""
    import torch
    g = np.load(os.path.join(GOLDEN, "dsscpp_reference_run.npz"))
    mg = np.load(os.path.join(GOLDEN, "etihad_mesh.npz"))
    m = A.MultiContext(etihad[0], A.default_params(), devices=list(range(torch.cuda.device_count())))
    mesh = A.Mesh(mg["vertices"], mg["triangles"])
    robot, sens, stages, gs = m.adaptive_sampling(mesh, g["grid_start"], g["grid_size"], start_res=3.0, res_decrement=1.5)
    assert len(stages) == 2 and stages["candidates"][0] == 2112 and stages["accepted"][0] == 335
    assert len(robot) == len(g["robot"]) == 2885
    assert np.array_equal(robot, g["robot"])
    assert np.array_equal(sens, g["sensors"])
    cand = np.zeros((4000, 7))
    n = A._abi.load().asscpp_grid_candidates(A._ptr(np.ascontiguousarray(g["grid_start"]), A.C.c_double),
                                             A._ptr(np.ascontiguousarray(g["grid_size"]), A.C.c_double), 3.0, 45.0,
                                             A._ptr(cand, A.C.c_double), len(cand))
    assert n == 2112
    mesh.close()
    m.close()


def test_view_cells_equal_oracle(A, oracle_mod, etihad_ctx, etihad_oracle, space_2p5):
    """ASSCPP_EVAL_VIEW_CELLS / asscpp_get_view_cells: per view, the coverage voxels its visible cloud touches (the
    occupied octomap keys of coverage_path_planning_heuristic.cpp:657-674 at volumetricVoxelRes 0.5) -- small-batch
    path, general path, grouped and not; the gains are unchanged by the flag."""
    o = etihad_oracle
    n_cov = etihad_ctx.n_cov_voxels
    for n, group in ((6, 1), (40, 2), (600, 1), (600, 2)):
        k = n // 2
        poses = np.stack([space_2p5["cam0"][600:600 + k], space_2p5["cam1"][600:600 + k]], 1).reshape(-1, 7)
        etihad_ctx.eval_ex(poses, group_size=group, flags=A.EVAL_VIEW_CELLS | A.EVAL_HASH)
        cells = etihad_ctx.view_cells(n)
        res = etihad_ctx.results(n)
        etihad_ctx.eval_ex(poses, group_size=group, flags=A.EVAL_HASH)
        plain = etihad_ctx.results(n)
        for f in ("visible", "gain_pts", "gain_vox", "vis_hash"):
            assert np.array_equal(res[f], plain[f]), (n, group, f)
        assert cells.shape == (n, n_cov)
        step = 1 if n <= 40 else 37
        for v in range(0, n, step):
            cp = np.zeros(o.pts_words(), np.uint64)
            cv = np.zeros(o.vox_words(), np.uint64)
            o.accumulate(poses[v:v + 1], cp, cv, min_visible=-1, nthreads=1)
            want = np.unpackbits(cv.view(np.uint8), bitorder="little")[:n_cov].astype(bool)
            assert np.array_equal(cells[v], want), (n, group, v)
        assert cells.any()
    with pytest.raises(A.AsscppError):
        etihad_ctx.eval_ex(np.tile(space_2p5["cam0"][600], (5000, 1)), flags=A.EVAL_VIEW_CELLS)


def test_view_cells_with_a_replayed_view(A, oracle_mod, monkeypatch):
    """ASSCPP_EVAL_VIEW_CELLS when a view of the batch takes the ordered replay (quirk Q5): the replay rewrites that
    view's row; the one-launch path falls back to the general one.  Scene of test_carry_inside_large_batch."""
    rng = np.random.default_rng(5)
    xyz = np.concatenate([np.array([(42, 6, 2), (14, 2, 2), (2.5, 2, 2), (14, 3, 6)], np.float32),
                          (rng.random((5000, 3)) * np.array([60, 30, 12]) + np.array([0, -40, 0])).astype(np.float32)])
    pa = A.default_params(leaf=4.0, far_dist=60.0, cov_leaf=2.0)
    po = oracle_mod.make_params(leaf=4.0, far=60.0, cov_leaf=2.0)
    poses = np.array([[1, 2, 2, 0, 0, 0, 1.0], [1, -20, 2, 0, 0, 0, 1.0], [5, -30, 6, 0, 0, 0.38268343, 0.92387953],
                      [1, 2, 2, 0, 0, 0, 1.0]])
    orc = oracle_mod.Oracle(xyz, po)
    want = []
    for v in range(len(poses)):
        cp = np.zeros(orc.pts_words(), np.uint64)
        cv = np.zeros(orc.vox_words(), np.uint64)
        orc.accumulate(poses[v:v + 1], cp, cv, min_visible=-1, nthreads=1)
        want.append(np.unpackbits(cv.view(np.uint8), bitorder="little")[:orc.n_cov].astype(bool))
    want = np.array(want)
    for no_small in (False, True):
        if no_small:
            monkeypatch.setenv("ASSCPP_NO_SMALL", "1")
        ctx = A.CoverageContext(xyz, pa)
        monkeypatch.delenv("ASSCPP_NO_SMALL", raising=False)
        for group in (1, 2):
            ctx.eval_ex(poses, group_size=group, flags=A.EVAL_VIEW_CELLS | A.EVAL_HASH)
            cells = ctx.view_cells(len(poses))
            res = ctx.results(len(poses))
            assert res["carry"][0] > 0 and (res["status"][0] & A.VS_ORDERED)
            assert np.array_equal(cells, want), (no_small, group)
        ctx.close()
    assert want.any()
