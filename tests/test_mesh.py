"""Mesh sample filters (coverage_path_planning_heuristic.cpp:300-335).

CPU: the oracle restatement against an independent numpy formulation (signed volumes for the crossings,
plane projection + edge distances for the stand-off).  GPU: the brute-force triangle kernel equals the oracle
bit for bit (crossing counts, squared distances as IEEE doubles, pass flags, free segments).
Parity unpinned: the reference answers these through CGAL's AABB tree, which is not in the image; CGAL's
predicates and this formulation can only differ for queries that graze an edge or a vertex, and the oracle's
grazing report bounds how close the fixture queries come to that.
"""
import os

import numpy as np
import pytest

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))


def etihad_mesh():
    g = np.load(os.path.join(HERE, "golden", "etihad_mesh.npz"))
    return g["vertices"], g["triangles"]


def box_mesh(lo, hi):
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    v = np.array([[x, y, z] for z in (lo[2], hi[2]) for y in (lo[1], hi[1]) for x in (lo[0], hi[0])], np.float32)
    q = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    t = np.array([[a, b, c] for a, b, c, d in q] + [[a, c, d] for a, b, c, d in q], np.int32)
    return v, t


def candidates(n, seed=0):
    """Sample positions around the aircraft the way the adaptive loop makes them (cluster boxes inflated)."""
    rng = np.random.default_rng(seed)
    return np.stack([rng.uniform(-22, 22, n), rng.uniform(-25, 25, n), rng.uniform(-3, 14, n)], 1)


# ---- independent numpy formulation ----
def np_crossings(tri, o, d, segment=False):
    p0, p1, p2 = tri[:, 0], tri[:, 1], tri[:, 2]

    def vol(a, b):  # signed volume of (o, o + d, a, b)
        return np.einsum("j,ij->i", d, np.cross(a - o, b - o))

    s0, s1, s2 = vol(p0, p1), vol(p1, p2), vol(p2, p0)
    same = ((s0 >= 0) & (s1 >= 0) & (s2 >= 0)) | ((s0 <= 0) & (s1 <= 0) & (s2 <= 0))
    nrm = np.cross(p1 - p0, p2 - p0)
    den = nrm @ d
    with np.errstate(divide="ignore", invalid="ignore"):
        t = np.einsum("ij,ij->i", nrm, p0 - o) / den
    ok = same & (den != 0) & (t >= 0)
    if segment:
        ok &= t <= 1
    margin = np.min(np.abs(np.stack([s0, s1, s2])), axis=0) / (np.abs(s0 + s1 + s2) + 1e-300)
    return int(ok.sum()), margin.min()


def np_sqdist(tri, p):
    a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]

    def seg(u, v):
        e = v - u
        t = np.clip(np.einsum("ij,ij->i", p - u, e) / np.einsum("ij,ij->i", e, e), 0, 1)
        r = p - (u + t[:, None] * e)
        return np.einsum("ij,ij->i", r, r)

    n = np.cross(b - a, c - a)
    nn = np.einsum("ij,ij->i", n, n)
    h = np.einsum("ij,ij->i", p - a, n)
    q = p - (h / nn)[:, None] * n
    inside = ((np.einsum("ij,ij->i", np.cross(b - a, q - a), n) >= 0)
              & (np.einsum("ij,ij->i", np.cross(c - b, q - b), n) >= 0)
              & (np.einsum("ij,ij->i", np.cross(a - c, q - c), n) >= 0))
    edge = np.minimum(np.minimum(seg(a, b), seg(b, c)), seg(c, a))
    return np.where(inside, h * h / nn, edge).min()


def test_oracle_box_inside_outside():
    v, t = box_mesh((-1, -2, -3), (2, 3, 4))
    m = O.Mesh(v, t)
    assert m.n_triangles == 12
    rng = np.random.default_rng(3)
    inside = rng.uniform((-0.9, -1.9, -2.9), (1.9, 2.9, 3.9), (200, 3))
    outside = inside + np.array([10.0, 0, 0])
    hits_in, sq_in, ok_in, _ = m.filter(inside, through=(100.0, 10.0, 56.0), min_dist=0.0, max_dist=100.0)
    hits_out, sq_out, ok_out, _ = m.filter(outside, through=(100.0, 10.0, 56.0), min_dist=0.0, max_dist=100.0)
    assert (hits_in % 2 == 1).all() and not ok_in.any()       # :322-325 odd -> inside -> rejected
    assert (hits_out % 2 == 0).all() and ok_out.all()
    # distance to an axis-aligned box has a closed form
    lo, hi = np.array([-1.0, -2, -3]), np.array([2.0, 3, 4])
    d_out = np.maximum(np.maximum(lo - outside, outside - hi), 0)
    np.testing.assert_allclose(sq_out, (d_out * d_out).sum(1), rtol=1e-13)
    d_in = np.minimum(inside - lo, hi - inside).min(1)
    np.testing.assert_allclose(sq_in, d_in * d_in, rtol=1e-12)
    # stand-off window :327-333 (both ends inclusive)
    p = np.array([[3.0, 0, 0], [4.0, 0, 0], [6.0, 0, 0], [6.5, 0, 0]])
    _, sq, ok, _ = m.filter(p, min_dist=1.0, max_dist=4.0)
    assert sq.tolist() == [1.0, 4.0, 16.0, 20.25] and ok.tolist() == [True, True, True, False]
    # segments :300-315
    a = np.array([[-5.0, 0, 0], [-5.0, 0, 0], [0.0, 0, 0], [5.0, 5, 5]])
    b = np.array([[5.0, 0.5, 0.2], [-2.0, 0.5, 0.2], [0.5, 0.5, 0.5], [5.0, -5, 5]])
    assert m.segments_free(a, b).tolist() == [False, True, True, True]


def test_oracle_mesh_against_numpy():
    v, t = etihad_mesh()
    m = O.Mesh(v, t)
    tri = v.astype(np.float64)[t]
    keep = np.linalg.norm(np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]), axis=1) != 0
    assert m.n_triangles == int(keep.sum())  # loadOBJFile skips collinear triangles
    tri = tri[keep]
    pos = candidates(150, seed=1)
    thr = np.array([100.0, 10.0, 56.0])
    hits, sqd, ok, graze = m.filter(pos, through=thr, nthreads=4)
    checked = 0
    for i in range(len(pos)):
        n, margin = np_crossings(tri, pos[i], thr - pos[i])
        np.testing.assert_allclose(sqd[i], np_sqdist(tri, pos[i]), rtol=1e-9)
        if min(margin, graze[i]) > 1e-9:  # both formulations agree away from edges and vertices
            assert hits[i] == n, i
            checked += 1
    assert checked > 100
    assert ((hits % 2 == 0) & (sqd >= 1.0) & (sqd <= 16.0) == ok).all()
    a, b = candidates(150, seed=2), candidates(150, seed=3)
    free = m.segments_free(a, b, nthreads=4)
    for i in range(len(a)):
        n, margin = np_crossings(tri, a[i], b[i] - a[i], segment=True)
        if margin > 1e-9:
            assert free[i] == (n == 0), i
    assert 0 < free.sum() < len(a)


@pytest.mark.gpu
def test_mesh_filters_equal_oracle():
    import asscpp_b200 as A

    v, t = etihad_mesh()
    om, gm = O.Mesh(v, t), A.Mesh(v, t)
    assert gm.n_triangles == om.n_triangles
    ss = np.load(os.path.join(HERE, "golden", "etihad_searchspace_2p5m.npz"))["uav"][:, :3].astype(np.float64)
    pos = np.concatenate([ss[:1500], candidates(1500, seed=5)])
    for thr, lo, hi in (((100.0, 10.0, 56.0), 1.0, 4.0), ((-30.0, 2.5, 80.0), 0.5, 2.0)):
        eh, es, eo, _ = om.filter(pos, through=thr, min_dist=lo, max_dist=hi, nthreads=16)
        gh, gs, go = gm.filter(pos, through=thr, min_dist=lo, max_dist=hi)
        assert np.array_equal(gh, eh)
        assert np.array_equal(gs.view(np.uint64), es.view(np.uint64))
        assert np.array_equal(go, eo)
        assert 0 < go.sum() < len(pos) and (gh % 2 == 1).any()
    a, b = candidates(3000, seed=6), candidates(3000, seed=7)
    assert np.array_equal(gm.segments_free(a, b), om.segments_free(a, b, nthreads=16))
    # empty batch, closed box
    h, s, o = gm.filter(np.zeros((0, 3)))
    assert len(h) == 0 and len(s) == 0 and len(o) == 0
    bv, bt = box_mesh((-1, -2, -3), (2, 3, 4))
    bm = A.Mesh(bv, bt)
    h, s, o = bm.filter(np.array([[0.0, 0, 0], [3.0, 0, 0]]), min_dist=1.0, max_dist=4.0)
    assert h.tolist() == [1, 0] and s.tolist() == [1.0, 1.0] and o.tolist() == [False, True]
    assert gm.L.asscpp_mesh_launch_count(gm.h) > 0
    gm.close(); bm.close()
