"""CPU tests (-m "not gpu"): the oracle against everything the reference ships that can pin it
(SURVEY.md section 4 / 8c), against hand-built micro-scenes, and against the committed golden vectors.

The reference has no tests and no recorded outputs for this path, so "pinning" is limited to
(1) the sensor-pose fixture tables (known answers of sspp/src/sensors.cpp:47-90),
(2) the accepted-waypoint table written by the reference's own ASSCPP run (weak known answer:
    every accepted waypoint saw more than 5 points, coverage_path_planning_heuristic.cpp:357),
(3) scenes small enough to derive the expected visible set by hand from the cited lines.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN

MOUNTS = (((0.0, 0.022, 0.065), (0.0, -0.349, 0.0)), ((0.0, 0.022, -0.065), (0.0, 0.349, 0.0)))  # coverage_heuristic_test.cpp:94-95


def _pose_close(p, q, pos_tol, quat_tol):
    a = np.asarray(p[3:])
    b = np.asarray(q[3:])
    if np.dot(a, b) < 0:
        a = -a  # <tf getRotation> sign is arbitrary
    return np.abs(np.asarray(p[:3]) - q[:3]).max() <= pos_tol and np.abs(a - b).max() <= quat_tol


@pytest.mark.parametrize("which", ["2p5", "1p5"])
def test_sensor_pose_known_answers(oracle_mod, space_2p5, space_1p5, which):
    """cscpp/txt/SearchSpaceUAV_* row k -> SearchSpaceCam_*_{0,1} row k (6 printed digits)."""
    ss = space_2p5 if which == "2p5" else space_1p5
    uav, cams = ss["uav"], (ss["cam0"], ss["cam1"])
    rows = range(len(uav)) if which == "2p5" else range(0, len(uav), 37)
    for k in rows:
        for cam, (off, rpy) in zip(cams, MOUNTS):
            s = oracle_mod.robot2sensor(uav[k], off, rpy)
            assert _pose_close(s, cam[k], 6e-5 * max(1.0, np.abs(cam[k, :3]).max() / 10 + 1), 6e-6), (k, s, cam[k])


def test_model_voxel_count_independent(oracle_mod, etihad):
    """OriginalVoxelsSize (occlusion_culling.cpp:50-65) recomputed with numpy: 2 759 on the bundled model."""
    xyz, vp = etihad
    inv = np.float32(1.0) / np.float32(0.5)
    ijk = np.floor(xyz * inv).astype(np.int64)
    n_np = len(np.unique(ijk, axis=0))
    o = oracle_mod.Oracle(xyz)
    assert o.n_cov == n_np == 2759
    assert o.n_occ == 2759
    assert np.allclose(vp[:3], 0.0)  # PCD VIEWPOINT translation = pcd_origin default


def test_accepted_waypoints_weak_known_answer(oracle_mod, etihad, space_2p5):
    """Every waypoint the reference's own ASSCPP run accepted must see > 5 points over its two sensors."""
    xyz, _ = etihad
    o = oracle_mod.Oracle(xyz)
    d = space_2p5["dsscpp"]  # all 3 357 complete rows
    poses = []
    for x, y, z, yaw in d:
        rp = np.array([x, y, z, 0, 0, np.sin(yaw / 2), np.cos(yaw / 2)])
        for off, rpy in MOUNTS:
            poses.append(oracle_mod.robot2sensor(rp, off, rpy))
    st = o.eval_batch(np.array(poses), memo=True, nthreads=os.cpu_count())
    vis = st[:, oracle_mod.ST_VISIBLE].reshape(-1, 2).sum(1)
    assert (vis > 5).all(), "accepted waypoints that see <= 5 points: %s" % np.nonzero(vis <= 5)[0]


def test_memo_equals_literal(oracle_mod, etihad, space_2p5):
    """Quirk Q1: one ray per occupied voxel gives the same sets as the reference's ray per point."""
    xyz, _ = etihad
    o = oracle_mod.Oracle(xyz)
    poses = np.concatenate([space_2p5["cam0"][500:700], space_2p5["cam1"][500:700]])
    a = o.eval_batch(poses, memo=True, nthreads=4)
    b = o.eval_batch(poses, memo=False, nthreads=4)
    cols = [c for c in range(oracle_mod.ST_COUNT) if c != oracle_mod.ST_STEPS]
    assert np.array_equal(a[:, cols], b[:, cols])
    assert a[:, oracle_mod.ST_VISIBLE].sum() > 10000


def test_golden_vectors(oracle_mod, etihad, space_2p5):
    """Committed outputs of the oracle on the bundled model (tools/make_golden.py): regression pin."""
    g = np.load(os.path.join(GOLDEN, "etihad_oracle_golden.npz"))
    xyz, _ = etihad
    o = oracle_mod.Oracle(xyz)
    poses = np.concatenate([space_2p5["cam0"], space_2p5["cam1"]])[g["view_index"]]
    st = o.eval_batch(poses, memo=True, nthreads=4)
    assert np.array_equal(st, g["stats"])
    r = o.eval_view(poses[int(g["list_view"])])
    assert np.array_equal(r["visible"], g["visible_list"])
    assert np.array_equal(r["survivors"], g["survivor_list"])


# ---- micro-scenes ------------------------------------------------------------------------------------
IDENT = np.array([0, 0, 0, 0, 0, 0, 1.0])  # camera at the origin, X forward, Y up, Z right


def _cells(i, js, ks, leaf=0.5):
    return [((i + 0.5) * leaf, (j + 0.5) * leaf, (k + 0.5) * leaf) for j in js for k in ks]


def test_wall_occludes_wall(oracle_mod):
    """A wall 3.25 m ahead hides a wall 5.25 m ahead (rays cross occupied cells farther than 2*leaf)."""
    front = _cells(6, range(-2, 2), range(-2, 2))
    back = _cells(10, range(-2, 2), range(-2, 2))
    xyz = np.array(front + back, np.float32)
    o = oracle_mod.Oracle(xyz)
    r = o.eval_view(IDENT)
    assert list(r["survivors"]) == list(range(32))
    assert list(r["visible"]) == list(range(16))
    assert r["stats"][oracle_mod.ST_CARRY] == 0


def test_self_occlusion_exemption(oracle_mod):
    """Quirk Q4: occupied cells within 2*leaf of the target never occlude (:558-562)."""
    xyz = np.array([(2.25, 0.25, 0.25), (2.75, 0.25, 0.25), (3.25, 0.25, 0.25), (3.75, 0.25, 0.25)], np.float32)
    r = oracle_mod.Oracle(xyz).eval_view(IDENT)
    assert list(r["survivors"]) == [0, 1, 2, 3]
    assert list(r["visible"]) == [0, 1, 2]  # the 4th cell is 1.5 m behind the 1st: occluded


def test_frustum_limits(oracle_mod):
    """near 0.7 / far 6.0 / 58 x 45 degrees (occlusion_culling.cpp:73-76), X forward."""
    pts = np.array([(0.5, 0, 0), (0.75, 0, 0), (5.9, 0, 0), (6.1, 0, 0), (3, 0, 1.6), (3, 0, 1.7), (3, 1.2, 0),
                    (3, 1.3, 0), (-3, 0, 0)], np.float32)
    r = oracle_mod.Oracle(pts).eval_view(IDENT)
    # tan(29 deg) * 3 = 1.663 (Z = right), tan(22.5 deg) * 3 = 1.243 (Y = up)
    assert list(r["survivors"]) == [1, 2, 4, 6]


def test_sticky_state_is_order_dependent(oracle_mod):
    """Quirk Q5: occlusionEstimation fails when tmin == -1.0f exactly and `state` keeps the previous
    point's value (occlusion_culling.cpp:308,332-335).  Same four points, three index orders."""
    p = oracle_mod.make_params(leaf=4.0, far=60.0, cov_leaf=4.0)
    pose = np.array([1, 2, 2, 0, 0, 0, 1.0])
    A, B, C, E = (42, 6, 2), (14, 2, 2), (2.5, 2, 2), (14, 3, 6)
    expect = {"ABCE": [3], "BCAE": [0, 1, 3], "EABC": [0]}
    for order, vis in expect.items():
        pts = np.array([{"A": A, "B": B, "C": C, "E": E}[ch] for ch in order], np.float32)
        o = oracle_mod.Oracle(pts, p)
        for memo in (False, True):
            r = o.eval_view(pose, memo=memo)
            assert list(r["visible"]) == vis, (order, memo)
            assert r["stats"][oracle_mod.ST_CARRY] == 2


def test_empty_and_blind(oracle_mod, etihad):
    o = oracle_mod.Oracle(np.zeros((0, 3), np.float32))
    r = o.eval_view(IDENT)
    assert len(r["survivors"]) == 0 and len(r["visible"]) == 0
    xyz, _ = etihad
    o = oracle_mod.Oracle(xyz)
    far_away = np.array([500.0, 500.0, 500.0, 0, 0, 0, 1.0])
    st = o.eval_batch(far_away[None, :])
    assert not st.any()


def test_planes_against_float64(oracle_mod, space_2p5):
    """frustum_culling.cpp:82-141 redone in float64 numpy: float32 planes agree to rounding."""
    o = oracle_mod.Oracle(np.zeros((1, 3), np.float32))
    for pose in space_2p5["cam0"][::97]:
        pl, T = o.planes(pose)
        x, y, z, w = pose[3:]
        s = 2.0 / (x * x + y * y + z * z + w * w)
        R = np.array([[1 - s * (y * y + z * z), s * (x * y - w * z), s * (x * z + w * y)],
                      [s * (x * y + w * z), 1 - s * (x * x + z * z), s * (y * z - w * x)],
                      [s * (x * z - w * y), s * (y * z + w * x), 1 - s * (x * x + y * y)]])
        view, up, right, t = R[:, 0], R[:, 1], R[:, 2], pose[:3]
        fh, fw = 2 * np.tan(np.radians(45) / 2) * 6.0, 2 * np.tan(np.radians(58) / 2) * 6.0
        fc = t + view * 6.0
        tl, tr, bl, br = fc + up * fh / 2 - right * fw / 2, fc + up * fh / 2 + right * fw / 2, \
            fc - up * fh / 2 - right * fw / 2, fc - up * fh / 2 + right * fw / 2
        a, b, c, d = bl - t, br - t, tr - t, tl - t
        ref = {0: np.cross(d, a), 1: np.cross(b, c), 2: np.cross(c, d), 3: np.cross(a, b),
               4: np.cross(bl - br, tr - br)}
        for k, nrm in ref.items():
            scale = np.abs(nrm).max()
            assert np.abs(pl[k, :3] - nrm).max() <= 2e-4 * scale, (k, pl[k], nrm)
        assert np.allclose(T, t.astype(np.float32))


def test_coverage_percent_float32(oracle_mod, etihad, space_2p5):
    """calcCoveragePercent (occlusion_culling.cpp:431-432): float(Matched)/float(Orig)*100 in float32."""
    xyz, _ = etihad
    o = oracle_mod.Oracle(xyz)
    r = o.eval_view(space_2p5["cam0"][610])
    ids = np.unique(o.cov_ids()[r["visible"]])
    want = np.float32(np.float32(len(ids)) / np.float32(2759)) * np.float32(100)
    assert np.float32(o.coverage_percent_idx(r["visible"])) == want
    assert len(ids) == r["stats"][oracle_mod.ST_GAIN_VOX]


def test_gain_against_covered_set(oracle_mod, etihad, space_2p5):
    xyz, _ = etihad
    o = oracle_mod.Oracle(xyz)
    poses = space_2p5["cam0"][600:640]
    cp = np.zeros(o.pts_words(), np.uint64)
    cv = np.zeros(o.vox_words(), np.uint64)
    o.accumulate(poses[:20], cp, cv, min_visible=5)
    st0 = o.eval_batch(poses)
    st1 = o.eval_batch(poses, covered_pts=cp, covered_vox=cv)
    big = st0[:20, oracle_mod.ST_VISIBLE] > 5
    assert (st1[:20][big][:, oracle_mod.ST_GAIN_PTS] == 0).all()
    assert (st1[:20][big][:, oracle_mod.ST_GAIN_VOX] == 0).all()
    assert (st1[:, oracle_mod.ST_GAIN_PTS] <= st0[:, oracle_mod.ST_GAIN_PTS]).all()
    # the union is a set: accumulating twice changes nothing
    cp2, cv2 = cp.copy(), cv.copy()
    o.accumulate(poses[:20], cp2, cv2, min_visible=5)
    assert np.array_equal(cp, cp2) and np.array_equal(cv, cv2)


def test_accuracy_terms_literal(oracle_mod, etihad, space_2p5):
    """calcAvgAccuracy (occlusion_culling.cpp:445-475) restated literally in numpy float64 on the oracle's
    own viewport transform equals orc_accuracy_idx bit for bit (same left-to-right double sum)."""
    O = oracle_mod
    xyz, _ = etihad
    o = O.Oracle(xyz)
    seen = 0
    for pose in space_2p5["cam1"][600:640:3]:
        vis = o.eval_view(pose)["visible"]
        if len(vis) == 0:
            continue
        seen += 1
        sd, sz, lo, hi = o.accuracy_idx(pose, vis)
        t = O.viewport_transform(pose, xyz[vis])
        ed = 0.0
        ez = 0.0
        for k in range(len(vis)):
            val = float(t[k, 0])
            ed += 0.0000285 * val * val
            vz = float(xyz[vis[k], 2])
            ez += 0.0000285 * vz * vz
        assert sd == ed and sz == ez
        assert lo == t[:, 0].min() and hi == t[:, 0].max()
    assert seen >= 3


def test_uncovered_clusters_against_scipy(oracle_mod, etihad, space_2p5):
    """pointCloudDiff + clusteringPointCloud + findClusterBB (heuristic.cpp:88-246) restated literally, pinned
    against an independent implementation: on a 0.5 m lattice a 0.8 m Euclidean tolerance links exactly the face
    and edge neighbours (0.5, 0.707 < 0.8 < 0.866), i.e. scipy's 18-connected labelling of the dense grid."""
    from scipy import ndimage
    O = oracle_mod
    xyz, _ = etihad
    o = O.Oracle(xyz)
    cv = np.zeros(o.vox_words(), np.uint64)
    cp = np.zeros(o.pts_words(), np.uint64)
    poses = np.concatenate([space_2p5["cam0"][::7], space_2p5["cam1"][::7]])
    o.accumulate(poses, cp, cv, min_visible=5, nthreads=4)
    ids, ctr = o.uncovered(cv)
    covered = int(np.unpackbits(cv.view(np.uint8)).sum())
    assert len(ids) == o.n_cov - covered and 0 < len(ids) < o.n_cov
    assert np.all(np.diff(ids) > 0)
    labels, k, ties = O.euclidean_clusters(ctr, 0.8, 5, 50000)
    assert ties == 0 and k > 1
    # independent: dense grid + 18-connectivity
    ijk = np.floor(ctr * np.float32(2.0)).astype(np.int64)
    lo = ijk.min(0)
    dense = np.zeros(ijk.max(0) - lo + 1, bool)
    dense[tuple((ijk - lo).T)] = True
    struct = ndimage.generate_binary_structure(3, 2)
    lab, nlab = ndimage.label(dense, structure=struct)
    mine = lab[tuple((ijk - lo).T)]
    sizes = np.bincount(mine)
    keep = (sizes[mine] >= 5) & (sizes[mine] <= 50000)
    assert np.array_equal(labels >= 0, keep)
    # same partition: one oracle label per scipy label and vice versa
    pairs = set(zip(labels[keep].tolist(), mine[keep].tolist()))
    assert len(pairs) == k == len(set(a for a, _ in pairs)) == len(set(b for _, b in pairs))
    # largest first
    csz = np.bincount(labels[labels >= 0])
    assert np.all(np.diff(csz) <= 0)
    # findClusterBB: centres of the extreme cells, +5 m in x, y and +3 m in z
    for c in range(min(k, 4)):
        pts = ctr[labels == c]
        start, size = O.cluster_bb(pts, 0.5)
        assert np.allclose(start, [pts[:, 0].min() - 2.5, pts[:, 1].min() - 2.5, pts[:, 2].min()])
        assert np.allclose(size, [np.ptp(pts[:, 0]) + 5, np.ptp(pts[:, 1]) + 5, np.ptp(pts[:, 2]) + 3])
