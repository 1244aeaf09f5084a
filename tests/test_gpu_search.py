"""GPU: the ORBmatcher searches outside the tracking path (relocalisation / Sim3 SearchByProjection, Fuse,
SearchBySim3, SearchForInitialization, SearchForTriangulation) through the C ABI against the CPU oracle —
identical indices, order-dependent bindings and tie cases included."""
import numpy as np
import pytest

from tests.search_cases import H, W, point_queries, sim3_case, triangulation_case, two_views

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tv():
    return two_views(21, nf=1500)


def _frame(k, d, sf, u_right=None):
    from orbslam3_cpp_fork_b200.matcher import Frame
    return Frame(k, d, sf, W, H, mvuRight=u_right, mb=0.1, mbf=40.0)


@pytest.mark.parametrize("seed,n,th,orb_dist,ori", [(1, 4000, 10.0, 100, True), (2, 2500, 15.0, 64, True), (3, 3000, 6.0, 90, False)])
def test_reloc_projection_vs_oracle(tv, seed, n, th, orb_dist, ori):
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ORBmatcher
    k, d, sf = tv["k1"], tv["d1"], tv["sf"]
    q = point_queries(k, d, sf, n, seed, jitter=6.0)
    F = _frame(k, d, sf)
    F.mvpMapPoints[q["bound"] > 0] = 10 ** 6                      # bound on entry
    m = ORBmatcher(0.9, ori)
    r, best, rot, keepb = m.SearchByProjectionReloc(F, q["desc"], q["u"], q["v"], q["level"], q["angle"], q["active"], th, orb_dist)
    fv, keep = op.make_frame_view(k, d, None, q["bound"], W, H, sf)
    rr, rbest, rrot, rkeep = op.search_by_projection_reloc(fv, q["desc"], q["u"], q["v"], q["level"], q["angle"], q["active"], th,
                                                           orb_dist, ori)
    assert r == rr and (best == rbest).all() and (rot == rrot).all() and (keepb == rkeep).all()
    hit = best[best >= 0]
    assert len(hit) > 300 and len(set(hit.tolist())) == len(hit) and (q["bound"][hit] == 0).all()


@pytest.mark.parametrize("seed,n,th,ratio", [(4, 4000, 10, 1.0), (5, 3000, 5, 1.5), (6, 2000, 3, 0.8)])
def test_sim3_projection_vs_oracle(tv, seed, n, th, ratio):
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ORBmatcher
    k, d, sf = tv["k2"], tv["d2"], tv["sf"]
    q = point_queries(k, d, sf, n, seed)
    m = ORBmatcher()
    r, best = m.SearchByProjectionSim3(_frame(k, d, sf), q["bound"], q["desc"], q["u"], q["v"], q["level"], q["active"], th, ratio)
    fv, keep = op.make_frame_view(k, d, None, q["bound"], W, H, sf)
    rr, rbest = op.search_by_projection_sim3(fv, q["desc"], q["u"], q["v"], q["level"], q["active"], th, ratio)
    assert r == rr and (best == rbest).all() and r > 100
    hit = best[best >= 0]
    assert len(set(hit.tolist())) == len(hit) and (q["bound"][hit] == 0).all()


@pytest.mark.parametrize("seed,n,th,check", [(7, 4000, 3.0, True), (8, 3000, 4.0, False), (9, 3000, 2.5, True)])
def test_fuse_vs_oracle(tv, seed, n, th, check):
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ORBmatcher
    k, d, sf = tv["k1"], tv["d1"], tv["sf"]
    q = point_queries(k, d, sf, n, seed, jitter=3.0)
    u_right = q["u_right"] if seed != 9 else None                 # seed 9: monocular KeyFrame (5.99 gate only)
    m = ORBmatcher()
    r, best = m.Fuse(_frame(k, d, sf, u_right), tv["inv_sigma2"] if check else None, q["desc"], q["u"], q["v"],
                     q["ur"] if check else None, q["level"], q["active"], th, check)
    fv, keep = op.make_frame_view(k, d, u_right, None, W, H, sf, mb=0.1, mbf=40.0)
    rr, rbest = op.fuse(fv, tv["inv_sigma2"] if check else None, q["desc"], q["u"], q["v"], q["ur"] if check else None,
                        q["level"], q["active"], th, check)
    assert r == rr and (best == rbest).all() and r > 100


@pytest.mark.parametrize("seed,th", [(10, 7.5), (11, 3.0)])
def test_search_by_sim3_vs_oracle(tv, seed, th):
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ORBmatcher
    a, b = sim3_case(tv, seed)
    sf = tv["sf"]
    m = ORBmatcher()
    r, m12 = m.SearchBySim3(_frame(tv["k1"], tv["d1"], sf), _frame(tv["k2"], tv["d2"], sf), *a, *b, th)
    fv1, k1 = op.make_frame_view(tv["k1"], tv["d1"], None, None, W, H, sf)
    fv2, k2 = op.make_frame_view(tv["k2"], tv["d2"], None, None, W, H, sf)
    rr, rm12 = op.search_by_sim3(fv1, fv2, *a, *b, th)
    assert r == rr and (m12 == rm12).all() and r > 100


@pytest.mark.parametrize("seed,tie_heavy,window,ratio,ori", [(12, False, 100, 0.9, True), (13, True, 60, 0.9, True),
                                                           (14, False, 30, 0.7, False), (15, True, 100, 1.0, True)])
def test_search_for_initialization_vs_oracle(seed, tie_heavy, window, ratio, ori):
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ORBmatcher
    t = two_views(seed, nf=2500, tie_heavy=tie_heavy)
    k1, d1, k2, d2, sf = t["k1"], t["d1"], t["k2"], t["d2"], t["sf"]
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    m = ORBmatcher(ratio, ori)
    mine = prev.copy()
    r, m12 = m.SearchForInitialization(_frame(k1, d1, sf), _frame(k2, d2, sf), mine, window)
    fv1, a = op.make_frame_view(k1, d1, None, None, W, H, sf); fv2, b = op.make_frame_view(k2, d2, None, None, W, H, sf)
    rr, rm12, rprev = op.search_for_initialization(fv1, fv2, prev, window, ratio, ori)
    assert r == rr and (m12 == rm12).all() and (mine == rprev).all() and r > 5
    # second call with the updated vbPrevMatched, as Tracking::MonocularInitialization does frame after frame
    r2, m12b = m.SearchForInitialization(_frame(k1, d1, sf), _frame(k2, d2, sf), mine, window)
    rr2, rm12b, _ = op.search_for_initialization(fv1, fv2, rprev, window, ratio, ori)
    assert r2 == rr2 and (m12b == rm12b).all()


@pytest.mark.parametrize("seed,only_stereo,coarse,ori", [(16, False, False, True), (17, True, False, True), (18, False, True, True),
                                                       (19, False, True, False)])
def test_search_for_triangulation_vs_oracle(tv, seed, only_stereo, coarse, ori):
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ORBmatcher
    from orbslam3_cpp_fork_b200.matcher import join_feature_vectors
    c = triangulation_case(tv, seed)
    sf = tv["sf"]
    m = ORBmatcher(0.6, ori)
    r, m12 = m.SearchForTriangulation(_frame(tv["k1"], tv["d1"], sf, c["ur1"]), _frame(tv["k2"], tv["d2"], sf, c["ur2"]),
                                      c["has1"], c["has2"], c["fv1"], c["fv2"], c["F12"], c["ep"], tv["sigma2"], only_stereo, coarse)
    fv1, a = op.make_frame_view(tv["k1"], tv["d1"], c["ur1"], c["has1"], W, H, sf)
    fv2, b = op.make_frame_view(tv["k2"], tv["d2"], c["ur2"], c["has2"], W, H, sf)
    off1, idx1, off2, idx2 = join_feature_vectors(c["fv1"], c["fv2"])
    rr, rm12 = op.search_for_triangulation(fv1, fv2, off1, idx1, off2, idx2, c["F12"], c["ep"], tv["sigma2"], only_stereo, coarse, ori)
    assert r == rr and (m12 == rm12).all() and r > (2 if only_stereo else 15)


def test_search_degenerate_inputs(tv):
    from orbslam3_cpp_fork_b200 import ORBmatcher
    k, d, sf = tv["k1"], tv["d1"], tv["sf"]
    m = ORBmatcher()
    q = point_queries(k, d, sf, 64, 1)
    none = np.zeros(64, np.uint8)
    r, best = m.SearchByProjectionSim3(_frame(k, d, sf), np.zeros(len(k), np.uint8), q["desc"], q["u"], q["v"], q["level"], none, 10, 1.0)
    assert r == 0 and (best == -1).all()                          # nothing active
    r, best = m.Fuse(_frame(k, d, sf), None, q["desc"], q["u"], q["v"], None, q["level"], q["active"], 3.0, False)
    everything = np.ones(len(k), np.uint8)
    r2, best2 = m.SearchByProjectionSim3(_frame(k, d, sf), everything, q["desc"], q["u"], q["v"], q["level"], q["active"], 10, 1.0)
    assert r2 == 0 and (best2 == -1).all()                        # every KeyFrame feature already matched
    e = _frame(k[:0], d[:0], sf)
    r3, m12 = m.SearchForInitialization(e, _frame(k, d, sf), np.zeros((0, 2), np.float32), 100)
    assert r3 == 0 and len(m12) == 0
