"""The C++ drop-in surface (include/cscpp/occlusion_culling_b200.h, include/sspp/coverage_heuristic_b200.h):
it compiles against the compat/ stand-ins (CPU test) and, on a GPU, reproduces the oracle's visible sets,
coverage percentages, SurfaceCoverageH walk and ASSCPP acceptance counts (gpu test)."""
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT

MOUNTS = (((0.0, 0.022, 0.065), (0.0, -0.349, 0.0)), ((0.0, 0.022, -0.065), (0.0, 0.349, 0.0)))


def build_shim(tmp_path, lib_built):
    exe = str(tmp_path / "test_shim")
    libdir = os.path.dirname(lib_built)
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(ROOT, "compat"),
                           "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "test_shim.cpp"),
                           "-L" + libdir, "-lasscpp_b200", "-Wl,-rpath," + libdir, "-o", exe])
    return exe


def test_shim_compiles_and_links(tmp_path, lib_built):
    exe = build_shim(tmp_path, lib_built)
    assert os.path.exists(exe)


def write_pcd(path, xyz):
    with open(path, "wb") as f:
        f.write(("# .PCD v0.7\nVERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nCOUNT 1 1 1\nWIDTH %d\nHEIGHT 1\n"
                 "VIEWPOINT 0 0 0 0 0.5 0 0\nPOINTS %d\nDATA binary\n" % (len(xyz), len(xyz))).encode())
        f.write(np.ascontiguousarray(xyz, "<f4").tobytes())


@pytest.mark.gpu
def test_shim_matches_oracle(tmp_path, lib_built, oracle_mod, etihad, space_2p5):
    O = oracle_mod
    xyz, _ = etihad
    exe = build_shim(tmp_path, lib_built)
    pcd = str(tmp_path / "model.pcd")
    write_pcd(pcd, xyz)
    rows = [600, 610, 633, 700, 742, 815, 900, 1010]
    poses = np.stack([space_2p5[c][k] for k in rows for c in ("cam0", "cam1")])
    np.savetxt(str(tmp_path / "poses.txt"), poses, fmt="%.17g")
    g = np.load(os.path.join(ROOT, "tests", "golden", "etihad_mesh.npz"))
    with open(str(tmp_path / "mesh.txt"), "w") as f:
        for v in g["vertices"]:
            f.write("v %.9g %.9g %.9g\n" % tuple(v))
        for t in g["triangles"]:
            f.write("f %d %d %d\n" % tuple(t))
    out = subprocess.check_output([exe, pcd, str(tmp_path / "poses.txt"), str(tmp_path / "mesh.txt")], text=True)
    o = O.Oracle(xyz)
    assert re.search(r"model %d points OriginalVoxelsSize 2759" % len(xyz), out)
    # 1. per-view visible set and calcCoveragePercent
    views = re.findall(r"view (\d+) visible (\d+) hash (\d+) cloud_matches (\d) coverage (\S+)", out)
    assert len(views) == len(poses)
    for k, nvis, h, same, pct in views:
        r = o.eval_view(poses[int(k)])
        assert int(nvis) == len(r["visible"]) and int(same) == 1
        assert int(h) == O.set_hash(r["visible"])
        assert np.float32(float(pct)) == np.float32(o.coverage_percent_idx(r["visible"]))
    # 2. SurfaceCoverageH walk (coverage_path_planning_heuristic.cpp:470-488) restated with the oracle
    ids = o.cov_ids()
    vox = [np.unique(ids[np.concatenate([o.eval_view(poses[2 * i])["visible"], o.eval_view(poses[2 * i + 1])["visible"]])])
           for i in range(len(rows))]
    pos = [poses[2 * i][:3] for i in range(len(rows))]

    def pct(s):
        return float(np.float32(np.float32(np.float32(len(s)) / np.float32(o.n_cov)) * np.float32(100)))

    covered = set(vox[0].tolist())
    cov, fval, cur = pct(covered), 0.0, 0  # root f = 0 (heuristic.cpp:1243-1247)
    m = re.search(r"root coverage (\S+) f (\S+)", out)
    assert float(m.group(1)) == pytest.approx(cov, abs=0) and float(m.group(2)) == pytest.approx(fval, abs=0)
    steps = re.findall(r"step (\d+) best (\d+) coverage (\S+) f (\S+)", out)
    assert len(steps) == 4
    for s, best, c_s, f_s in steps:
        cand = []
        for i in range(1, len(rows)):
            u = covered | set(vox[i].tolist())
            c = pct(u)
            d = float(np.linalg.norm(pos[i] - pos[cur]))
            gain = c - cov
            cand.append((fval + ((1 / d) * gain if d != 0 else gain), i, c, u))
        bi = max(range(len(cand)), key=lambda j: (cand[j][0], -j))
        fbest, ibest, cbest, ubest = cand[bi]
        assert int(best) == ibest, (s, best, ibest)
        assert float(c_s) == pytest.approx(cbest, abs=1e-6) and float(f_s) == pytest.approx(fbest, rel=1e-12, abs=1e-9)
        covered, cov, fval, cur = ubest, cbest, fbest, ibest
    assert int(re.search(r"walk launches (\d+)", out).group(1)) == 4  # one scoring launch per expansion
    # 2b. default mode = what the reference's A* computes (quirk Q10): each child against the EMPTY set
    q10 = re.findall(r"q10 (\d+) coverage (\S+) f (\S+)", out)
    assert len(q10) == len(rows) - 1
    root_cov = pct(set(vox[0].tolist()))
    for i_s, c_s, f_s in q10:
        i = int(i_s)
        c = pct(set(vox[i].tolist()))
        d = float(np.linalg.norm(pos[i] - pos[0]))
        gain = c - root_cov
        assert float(c_s) == pytest.approx(c, abs=0)
        assert float(f_s) == pytest.approx((1 / d) * gain if d != 0 else gain, rel=1e-12, abs=1e-12)
    assert int(re.search(r"q10 launches (\d+)", out).group(1)) == 1
    # 3. acceptance rule (:339-357) and uncovered percentage (:143)
    st = o.eval_batch(poses, memo=True)
    vis = st[:, O.ST_VISIBLE].reshape(-1, 2)
    cv = np.zeros(o.vox_words(), np.uint64)
    cp = np.zeros(o.pts_words(), np.uint64)
    o.accumulate(poses, cp, cv, min_visible=5)
    nvox = int(np.unpackbits(cv.view(np.uint8)).sum())
    m = re.search(r"accepted (\d+) of (\d+) uncovered (\S+)", out)
    assert int(m.group(1)) == int((vis.sum(1) > 5).sum()) and int(m.group(2)) == len(rows)
    assert float(m.group(3)) == pytest.approx((o.n_cov - nvox) / o.n_cov * 100, rel=1e-12)
    ids_u, ctr_u = o.uncovered(cv)
    lab_u, k_u, _ = O.euclidean_clusters(ctr_u, 0.8, 5, 50000)
    m = re.search(r"clusters (\d+)(.*)", out)
    assert int(m.group(1)) == k_u
    for c, chunk in enumerate(x for x in m.group(2).split("|") if x.strip()):
        t = chunk.split()
        start, size = O.cluster_bb(ctr_u[lab_u == c], 0.5)
        assert int(t[0]) == int((lab_u == c).sum())
        assert [float(x) for x in t[1:4]] == start.tolist() and [float(x) for x in t[4:7]] == size.tolist()
    # 4. accuracy terms (occlusion_culling.cpp:445-520): the shim's host loops are the reference's loops on the same
    #    float32 depths, so the averages are bit-identical; SSMaxMinAccuracy comes from the batched GPU extrema
    accs = dict((int(k), (float(a), float(b))) for k, a, b in re.findall(r"accuracy (\d+) pose (\S+) z (\S+)", out))
    dmax, dmin = np.finfo(np.float64).tiny, np.finfo(np.float64).max
    nseen = 0
    for k, pose in enumerate(poses):
        vis_k = o.eval_view(pose)["visible"]
        if len(vis_k) == 0:
            assert k not in accs
            continue
        nseen += 1
        sd, sz, lo, hi = o.accuracy_idx(pose, vis_k)
        assert accs[k] == (sd / len(vis_k), sz / len(vis_k))
        dmax, dmin = max(dmax, float(hi)), min(dmin, float(lo))
    assert nseen > 0
    m = re.search(r"ssmaxmin max (\S+) min (\S+) set 1", out)
    assert float(m.group(1)) == 0.0000285 * dmax * dmax and float(m.group(2)) == 0.0000285 * dmin * dmin
    # 5. SurfaceCoveragewithAccuracyH (heuristic.cpp:490-513): f = parent.f + (1/d) * c * a with
    #    a = (maxAccuracyError - avg(0.0000285 z^2 over both sensors' visible clouds)) / maxAccuracyError.
    #    The average comes from exact GPU sums; the reference's running double sum may differ by n/2 ulp.
    max_err = 0.0000285 * dmax * dmax
    m = re.search(r"acc_root coverage (\S+) f (\S+) maxDepth (\S+)", out)
    cov0 = pct(set(vox[0].tolist()))
    assert float(m.group(1)) == cov0 and float(m.group(2)) == 0.0
    assert float(m.group(3)) == pytest.approx(np.sqrt(max_err / 0.0000285), rel=1e-15)
    kids = re.findall(r"acc_child (\d+) coverage (\S+) f (\S+)", out)
    assert len(kids) == len(rows) - 1
    for i, c_s, f_s in kids:
        i = int(i)
        both = np.concatenate([o.eval_view(poses[2 * i])["visible"], o.eval_view(poses[2 * i + 1])["visible"]])
        cov_i = pct(set(vox[0].tolist()) | set(vox[i].tolist()))
        assert float(c_s) == cov_i
        d = float(np.linalg.norm(pos[i] - pos[0]))
        if len(both) == 0:
            assert np.isnan(float(f_s))
            continue
        z = xyz[both, 2].astype(np.float64)
        avg = float(np.sum(0.0000285 * z * z)) / len(both)
        a = (max_err - avg) / max_err
        assert float(f_s) == pytest.approx((1 / d) * (cov_i - cov0) * a, rel=1e-12, abs=1e-12)
    # 7. VolumetricCoverageH (heuristic.cpp:581-641): f = parent.f + (accW a + distW normDist + angleW normAngle) * extra voxels
    m = re.search(r"vol_root coverage (\S+) f (\S+)", out)
    assert float(m.group(1)) == len(vox[0]) / o.n_cov * 100 and float(m.group(2)) == 0.0
    vkids = re.findall(r"vol_child (\d+) coverage (\S+) f (\S+)", out)
    assert len(vkids) == len(rows) - 1
    for i, c_s, f_s in vkids:
        i = int(i)
        both = np.concatenate([o.eval_view(poses[2 * i])["visible"], o.eval_view(poses[2 * i + 1])["visible"]])
        union = set(vox[0].tolist()) | set(vox[i].tolist())
        assert float(c_s) == len(union) / o.n_cov * 100
        extra = float(len(union) - len(vox[0]))
        d = float(np.linalg.norm(pos[i] - pos[0]))
        if len(both) == 0:
            assert np.isnan(float(f_s))
            continue
        z = xyz[both, 2].astype(np.float64)
        a = (max_err - float(np.sum(0.0000285 * z * z)) / len(both)) / max_err
        qa, qb = poses[0][3:], poses[2 * i][3:]
        dt = float(qa @ qb)
        ang = 2 * np.arccos(abs(dt) / np.sqrt(float(qa @ qa) * float(qb @ qb)))
        norm_angle = 1 - ang / (2 * np.pi)
        want = ((0.8 * a) + (0.5 * (2.5 - d) / 2.5) + (0.2 * norm_angle)) * extra if d != 0 else \
            ((0.8 * a) + (0.2 * norm_angle)) * extra
        assert float(f_s) == pytest.approx(want, rel=1e-11, abs=1e-11)
    # 6. mesh sample filters (heuristic.cpp:300-335)
    om = O.Mesh(g["vertices"], g["triangles"])
    nxt = np.concatenate([poses[2:], poses[:2]])
    assert [int(x) for x in re.search(r"mesh pass(.*)", out).group(1).split()] == \
        om.filter(poses[:, :3], (100.0, 10.0, 56.0), 1.0, 4.0)[2].astype(int).tolist()
    free = om.segments_free(poses[:, :3], nxt[:, :3]).astype(int).tolist()
    assert [int(x) for x in re.search(r"mesh free(.*)", out).group(1).split()] == free
    assert int(re.search(r"mesh single (\d)", out).group(1)) == free[0]
