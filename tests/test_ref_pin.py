"""Pins the oracle to the REFERENCE'S OWN CODE: oracle/_ref/libasscpp_ref.so is cscpp/src/{frustum_culling,
voxel_grid_occlusion_estimation, occlusion_culling, coverage_path_planning_heuristic}.cpp and sspp/src/*.cpp compiled
from /root/reference (oracle/Makefile) against the stand-in third-party headers under compat/.  Every comparison here
is reference object vs oracle restatement on the same inputs, bit for bit.

The library is built where /root/reference is mounted (this container) and shipped prebuilt; where neither the library
nor the reference exists the tests skip.
"""
import os

import numpy as np
import pytest

from conftest import ROOT

NTHREADS = os.cpu_count() or 1
REF_PCD = "/root/reference/cscpp/pcd/etihad_nowheels_nointernal_scaled_newdensed.pcd"


@pytest.fixture(scope="module")
def R():
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref is not built and /root/reference is not mounted")
    ref.build()
    return ref


@pytest.fixture(scope="module")
def ref_oc(R, etihad):
    # from the model points (the PointCloud::Ptr constructor, occlusion_culling.cpp:81-135): works on the GPU box too
    oc = R.RefOcclusionCulling(xyz=etihad[0], nthreads=NTHREADS)
    yield oc
    oc.close()


def test_constructor_counts_model_voxels(R, ref_oc, oracle_mod, etihad):
    """OriginalVoxelsSize (occlusion_culling.cpp:50-65): 2 759 on the bundled model, both constructors."""
    assert ref_oc.original_voxels == 2759.0
    assert oracle_mod.Oracle(etihad[0]).n_cov == 2759
    if os.path.exists(REF_PCD):
        oc = R.RefOcclusionCulling(pcd_path=REF_PCD)
        assert oc.n == len(etihad[0]) and oc.original_voxels == 2759.0
        oc.close()


def test_visible_and_survivor_lists_verbatim(ref_oc, oracle_mod, etihad, space_2p5):
    """extractVisibleSurface (occlusion_culling.cpp:265-373): FrustumCloud and the returned cloud, as index lists."""
    o = oracle_mod.Oracle(etihad[0])
    checked = 0
    for k in list(range(0, 1225, 9)):
        for cam in ("cam0", "cam1"):
            s, v = ref_oc.extract(space_2p5[cam][k])
            r = o.eval_view(space_2p5[cam][k], memo=False)
            assert np.array_equal(s, r["survivors"]), (k, cam)
            assert np.array_equal(v, r["visible"]), (k, cam)
            checked += len(v)
    assert checked > 50000


@pytest.mark.parametrize("which", ["2p5", "1p5"])
def test_every_fixture_view_counts_and_set_hashes(ref_oc, oracle_mod, etihad, space_2p5, space_1p5, which):
    """All 2 x 1 225 (cscpp/txt) and 2 x 17 583 (sspp/resources) sensor poses of config 1: survivor count, visible
    count and the 64-bit hash of the visible index set, reference object vs oracle (literal per-point loop)."""
    sp = space_2p5 if which == "2p5" else space_1p5
    poses = np.concatenate([sp["cam0"], sp["cam1"]])
    st_ref = ref_oc.eval_batch(poses)
    o = oracle_mod.Oracle(etihad[0])
    st = o.eval_batch(poses, memo=False, nthreads=NTHREADS)
    assert np.array_equal(st_ref[:, 0], st[:, oracle_mod.ST_SURVIVORS])
    assert np.array_equal(st_ref[:, 1], st[:, oracle_mod.ST_VISIBLE])
    assert np.array_equal(st_ref[:, 2].astype(np.uint64), st[:, oracle_mod.ST_HASH].astype(np.uint64))
    assert int(st_ref[:, 1].sum()) > 100000


def test_coverage_percent_and_accuracy(ref_oc, oracle_mod, etihad, space_2p5):
    """calcCoveragePercent (:375-444) float32-equal; calcAvgAccuracy both overloads (:445-475) double-equal (same
    left-to-right sums); pointCloudViewportTransform (:538-618) float32-equal."""
    o = oracle_mod.Oracle(etihad[0])
    xyz = etihad[0]
    for k in (600, 610, 633, 700, 815):
        pose = space_2p5["cam0"][k]
        vis = o.eval_view(pose)["visible"]
        if len(vis) == 0:
            continue
        assert np.float32(ref_oc.coverage_percent_idx(vis)) == np.float32(o.coverage_percent_idx(vis))
        sd, sz, lo, hi = o.accuracy_idx(pose, vis)
        a_pose, a_z = ref_oc.avg_accuracy_idx(pose, vis)
        assert a_pose == sd / len(vis) and a_z == sz / len(vis)
        t_ref = ref_oc.viewport_transform(pose, xyz[vis])
        t_orc = oracle_mod.viewport_transform(pose, xyz[vis])
        assert np.array_equal(t_ref, t_orc)


def test_robot2sensor(R, oracle_mod, space_2p5):
    """Sensors::robot2sensorTransformation (sspp/src/sensors.cpp:47-90), the demo's two mounts."""
    mounts = (((0.0, 0.022, 0.065), (0.0, -0.349, 0.0)), ((0.0, 0.022, -0.065), (0.0, 0.349, 0.0)))
    for k in range(0, 1225, 37):
        for off, rpy in mounts:
            # the reference keeps the mount in float32 (Eigen::Vector3f, sensors.h:60); hand the oracle the same values
            off32 = np.asarray(off, np.float32).astype(np.float64)
            rpy32 = np.asarray(rpy, np.float32).astype(np.float64)
            a = R.robot2sensor(space_2p5["uav"][k], off, rpy)
            b = oracle_mod.robot2sensor(space_2p5["uav"][k], off32, rpy32)
            assert np.array_equal(a, b), (k, a, b)


def test_synthetic_fine_grid(R, oracle_mod):
    """leaf 0.05 m on a synthetic structure (the grid of configs 3 and 5): voxelRes is a public member the per-view grid
    reads (occlusion_culling.cpp:305)."""
    import ctypes
    L = ctypes.CDLL(os.path.join(ROOT, "asscpp_b200", "lib", "libasscpp_host.so"))
    L.asscpp_synth_structure_v1.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_void_p]
    L.asscpp_synth_views_v1.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_int64, ctypes.c_void_p]
    xyz = np.empty((300000, 3), np.float32)
    poses = np.empty((48, 7), np.float64)
    assert L.asscpp_synth_structure_v1(0xA55C0001, 0xA55C0002, len(xyz), xyz.ctypes.data) == 0
    assert L.asscpp_synth_views_v1(0xA55C0001, 0xA55C0003, len(poses), poses.ctypes.data) == 0
    for leaf in (0.05, 0.2):
        oc = R.RefOcclusionCulling(xyz=xyz, leaf=leaf, nthreads=min(NTHREADS, 8))
        st_ref = oc.eval_batch(poses)
        oc.close()
        st = oracle_mod.Oracle(xyz, oracle_mod.make_params(leaf=leaf)).eval_batch(poses, memo=False, nthreads=NTHREADS)
        assert np.array_equal(st_ref[:, 0], st[:, oracle_mod.ST_SURVIVORS])
        assert np.array_equal(st_ref[:, 1], st[:, oracle_mod.ST_VISIBLE])
        assert np.array_equal(st_ref[:, 2].astype(np.uint64), st[:, oracle_mod.ST_HASH].astype(np.uint64))
        assert int(st_ref[:, 1].sum()) > 1000


def test_first_sampling_stage_of_the_shipped_run_is_reproduced(R):
    """The one output of the reference's own run that ships with it (cscpp/txt/..._dsscpp.txt, written at
    coverage_heuristic_test.cpp:167-181): its first sampling stage -- generateRegularGrid + mesh filters + "sensors see
    more than 5 points" (pathplanner.cpp:106-201, heuristic.cpp:317-364) -- is reproduced row for row by the reference's
    own planner code running on the compiled reference; the only rows the run adds are the earliest ones, which the
    file lost where it is truncated.  tools/reproduce_dsscpp.py has the derivation of the grid and the full report."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import reproduce_dsscpp as D
    if not os.path.exists(D.FILE):
        pytest.skip("/root/reference is not mounted")
    rows = D.load_file()
    k = D.first_stage_block(rows)
    assert len(rows) - k == 302
    size = [D.UPPER[c] - D.LOWER[c] + 1.0 for c in range(3)]
    start = [D.LOWER[c] + size[c] / 2 for c in range(3)]
    p = R.RefPlanner(D.OBJ, D.PCD, heuristic_type=0)
    p.regular_grid(start, size, 3.0)
    got, _, _ = p.search_space()
    p.close()
    have = set(D.key(r) for r in got)
    missing = [r for r in rows[k:] if D.key(r) not in have]
    assert not missing, "rows of the file the run does not accept: %s" % missing[:5]
    before, after = D.extras_vs_truncation(got, rows)
    assert after == 0 and before == len(got) - 302


def test_uncovered_regions_of_the_adaptive_loop(R, oracle_mod, etihad):
    """pointCloudDiff (heuristic.cpp:88-145) incl. its unchecked linear-index lookup (quirk Q11),
    clusteringPointCloud (:147-190) and findClusterBB (:218-246): the reference's heuristic object vs the oracle, on the
    covered cloud the first sampling stage of the shipped run leaves behind."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import reproduce_dsscpp as D
    if not os.path.exists(D.OBJ):
        pytest.skip("/root/reference is not mounted")
    xyz = etihad[0]
    g = np.load(os.path.join(ROOT, "tests", "golden", "dsscpp_reference_run.npz"))
    o = oracle_mod.Oracle(xyz)
    stage1 = g["sensors"][-335:].reshape(-1, 7)
    cp = np.zeros(o.pts_words(), np.uint64)
    cv = np.zeros(o.vox_words(), np.uint64)
    o.accumulate(stage1, cp, cv, min_visible=5, nthreads=NTHREADS)
    idx = np.nonzero(np.unpackbits(cp.view(np.uint8), bitorder="little")[:len(xyz)])[0]
    p = R.RefPlanner(D.OBJ, D.PCD, 0)
    pct, ctr = p.point_cloud_diff(xyz[idx])
    ids_a, ctr_a = o.uncovered(cv, pcl_aliasing=True)
    ids_p, ctr_p = o.uncovered(cv)
    assert np.array_equal(ctr, ctr_a)                      # same cells, same order, same float32 centres
    assert pct == len(ctr_a) / o.n_cov * 100
    assert len(ctr_p) > len(ctr_a)                         # the quirk hides cells: plain set difference finds more
    labels, boxes = p.clusters(ctr)
    lab_o, k_o, _ = oracle_mod.euclidean_clusters(ctr, 0.8, 5, 50000)
    assert len(boxes) == k_o and np.array_equal(labels, lab_o)
    for c in range(k_o):
        start, size = oracle_mod.cluster_bb(ctr[lab_o == c], 0.5)
        assert boxes[c].tolist() == start.tolist() + size.tolist()
    p.close()


IG_CHAIN = [600, 601, 610, 633, 700, 815, 816]


def info_gain_chain_restated(oracle_mod, xyz, sp, chain, cells_of_view=None):
    """InfoGainVolumetricH (heuristic.cpp:643-759, 974-1085, 1128-1225) restated on top of the oracle's visible sets, in the
    form in which the block is well defined (a child's parent is the real scored node).  A sensor's occupied octomap keys
    = the 0.5 m cells floor(2 x) of its visible points (<octomap coordToKey>), keyToCoord = (cell + 0.5) * 0.5, the tree a
    dict of float32 log-odds, setNodeValue clamped to [logodds(0.5), logodds(0.9999999)].  Keys are walked in ascending
    order (the reference walks an unordered set: same terms, another order of the double sums).
    include/cscpp/info_gain_b200.h is this function in C++ with the visible cells coming from the GPU."""
    import math
    o = oracle_mod.Oracle(xyz)
    f32 = np.float32
    logodds = lambda p: f32(math.log(p / (1 - p)))
    prob_of = lambda l: 1. - (1. / (1. + math.exp(float(l))))
    # SSMaxMinAccuracy (occlusion_culling.cpp:477-520) over the whole search space
    mx = 2.2250738585072014e-308
    for cam in ("cam0", "cam1"):
        for pose in sp[cam]:
            vis = o.eval_view(pose)["visible"]
            if len(vis):
                mx = max(mx, float(o.accuracy_idx(pose, vis)[3]))
    max_err = 0.0000285 * mx * mx
    max_depth = math.sqrt(max_err / 0.0000285)
    cells_all = np.unique(np.floor(xyz.astype(np.float64) * 2.0).astype(np.int64), axis=0)
    model_volume = len(cells_all) * (0.5 * 0.5 * 0.5)
    model_entropy = len(cells_all) * (-1 * (0.5) * (math.log(0.5) / math.log(2)))
    lo, hi = logodds(0.5), logodds(0.9999999)
    rows, prev, tree = [], None, {}
    for i, k in enumerate(chain):
        tree = dict(tree)
        total = prev["total"] if prev else model_entropy
        vol = prev["vol"] if prev else 0.0
        nvox = prev["nvox"] if prev else 0
        for cam in ("cam0", "cam1"):
            pose = sp[cam][k]
            if cells_of_view is not None:
                cells = cells_of_view(pose)
            else:
                vis = o.eval_view(pose)["visible"]
                cells = np.unique(np.floor(xyz[vis].astype(np.float64) * 2.0).astype(np.int64), axis=0) if len(vis) else np.zeros((0, 3), np.int64)
            if len(cells) == 0:
                continue
            centres = ((cells.astype(np.float64) + 0.5) * 0.5).astype(np.float32)
            depth = oracle_mod.viewport_transform(pose, centres)[:, 0]
            # ascending compact id = ascending (k, j, i): the order info_gain_b200.h walks
            order = np.lexsort((cells[:, 0], cells[:, 1], cells[:, 2]))
            for j in order:
                key = tuple(cells[j])
                dist = min(float(depth[j]), max_depth) if float(depth[j]) > max_depth else float(depth[j])
                norm_acc = (max_err - (0.0000285 * dist * dist * 0.5)) / max_err
                if key in tree:
                    prob = prob_of(tree[key])
                    before = (-1 * prob * (math.log(prob) / math.log(2)))
                    post = (norm_acc * prob) / ((norm_acc * prob) + ((1 - norm_acc) * (1 - prob)))
                else:
                    nvox += 1
                    before = (-1 * (0.5) * (math.log(0.5) / math.log(2)))
                    post = (norm_acc * 0.5) / ((norm_acc * 0.5) + ((1 - norm_acc) * (1 - 0.5)))
                    vol += 0.5 * 0.5 * 0.5
                total -= before
                tree[key] = min(max(logodds(post), lo), hi)
                total += (-1 * post * (math.log(post) / math.log(2)))
        cov = (vol / model_volume) * 100
        if prev:
            d = float(np.sqrt(((sp["uav"][k][:3] - sp["uav"][chain[i - 1]][:3]) ** 2).sum()))
            dist_total = prev["dist"] + d
            f = prev["f"] + (total - prev["total"]) * math.exp(d * -0.2)
        else:
            dist_total, f = 0.0, total
        prev = {"total": total, "vol": vol, "nvox": nvox, "f": f, "dist": dist_total}
        rows.append([total, vol, cov, f, dist_total, nvox])
    return np.array(rows)


def test_info_gain_block_on_real_parents_restated(R, oracle_mod, etihad, space_2p5, tmp_path):
    """The reference's own InfoGainVolumetricH block (compiled from /root/reference, against the stand-in octomap of
    compat/) on a chain of nodes with real parents vs the restatement above: counts and volumes equal, entropies and f
    equal to the order of the sums."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import reproduce_dsscpp as D
    if not os.path.exists(D.OBJ):
        pytest.skip("/root/reference is not mounted")
    files = []
    for key in ("uav", "cam0", "cam1"):
        f = str(tmp_path / (key + ".txt"))
        np.savetxt(f, space_2p5[key], fmt="%.17g")
        files.append(f)
    p = R.RefPlanner(D.OBJ, D.PCD, heuristic_type=5, target=20.0, tolerance=1.0, conn_radius=2.5)
    p.load_grid(*files)
    p.set_accuracy_range()
    sens = np.stack([space_2p5["cam0"][IG_CHAIN], space_2p5["cam1"][IG_CHAIN]], 1)
    ref_rows = p.infogain_chain(space_2p5["uav"][IG_CHAIN], sens)
    p.close()
    mine = info_gain_chain_restated(oracle_mod, etihad[0], space_2p5, IG_CHAIN)
    assert np.array_equal(ref_rows[:, [1, 5]], mine[:, [1, 5]])            # coveredVolume, coveredVoxelsNum
    assert np.array_equal(ref_rows[:, 2], mine[:, 2]) and np.array_equal(ref_rows[:, 4], mine[:, 4])  # coverage, distance
    assert np.allclose(ref_rows[:, 0], mine[:, 0], rtol=1e-13, atol=0) and np.allclose(ref_rows[:, 3], mine[:, 3], rtol=1e-13, atol=0)
    assert ref_rows[-1, 5] > ref_rows[0, 5] > 0 and ref_rows[-1, 0] < ref_rows[0, 0]
