"""GPU (-m gpu): the ORBmatcher searches outside the tracking path through the C ABI against the
reference's OWN functions — the whole unmodified src/ORBmatcher.cc in oracle/_ref/liborb_ref.so, driven through
oracle/ref_compat/ref_capi_search.cpp (built in the container, shipped prebuilt to the GPU box).  Bit-exact on every
match index and count.  tests/ref_search_cases.py holds how the flat inputs of libcuorb relate to the KeyFrame /
MapPoint objects the reference is given; tests/test_oracle_vs_ref_search.py is the same comparison for the oracle."""
import numpy as np
import pytest

from oracle import oracle_py as op
from oracle import ref_py as rp
from tests.bow_cases import make_pair
from tests.ref_search_cases import (binding_from_best, bow_views, fuse_depths, has_mp_codes, in_image,
                                    triangulation_poses)
from tests.search_cases import H, W, point_queries, sim3_case, triangulation_case, two_views

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not rp.available(), reason="oracle/_ref/liborb_ref.so did not travel to this box")]


@pytest.fixture(scope="module")
def tv():
    return two_views(21, nf=1500)


def _frame(k, d, sf, u_right=None):
    from orbslam3_cpp_fork_b200.matcher import Frame
    return Frame(k, d, sf, W, H, mvuRight=u_right, mb=0.1, mbf=40.0)


@pytest.mark.parametrize("seed,tie_heavy,ratio,ori", [(1, False, 0.7, True), (2, True, 0.9, True), (3, False, 0.6, False)])
def test_search_by_bow_vs_reference(seed, tie_heavy, ratio, ori):
    from orbslam3_cpp_fork_b200 import ORBmatcher
    c = make_pair(seed, tie_heavy=tie_heavy)
    fv1, fv2, keep = bow_views(c)
    h1 = has_mp_codes(c["valid1"], seed); h2 = has_mp_codes(c["elig2"], seed + 1)
    m = ORBmatcher(ratio, ori)
    n, match = m.SearchByBoW_KF_F(c["desc1"], h1 == 1, c["angle1"], c["fv1"], c["desc2"], c["angle2"], c["fv2"])
    rn, rmatch = rp.search_by_bow(0, fv1, h1, c["fv1"], fv2, None, c["fv2"], ratio, ori)
    assert n == rn and (match == rmatch).all() and n > 50
    n, match = m.SearchByBoW_KF_KF(c["desc1"], h1 == 1, c["angle1"], c["fv1"], c["desc2"], h2 == 1, c["angle2"], c["fv2"])
    rn, rmatch = rp.search_by_bow(1, fv1, h1, c["fv1"], fv2, h2, c["fv2"], ratio, ori)
    assert n == rn and (match == rmatch).all() and n > 30


@pytest.mark.parametrize("seed,tie_heavy,window,ratio,ori", [(12, False, 100, 0.9, True), (13, True, 60, 0.9, True),
                                                           (15, True, 100, 1.0, False)])
def test_search_for_initialization_vs_reference(seed, tie_heavy, window, ratio, ori):
    from orbslam3_cpp_fork_b200 import ORBmatcher
    t = two_views(seed, nf=2500, tie_heavy=tie_heavy)
    k1, d1, k2, d2, sf = t["k1"], t["d1"], t["k2"], t["d2"], t["sf"]
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    m = ORBmatcher(ratio, ori)
    mine = prev.copy()
    r, m12 = m.SearchForInitialization(_frame(k1, d1, sf), _frame(k2, d2, sf), mine, window)
    fv1, a = op.make_frame_view(k1, d1, None, None, W, H, sf); fv2, b = op.make_frame_view(k2, d2, None, None, W, H, sf)
    rr, rm12, rprev = rp.search_for_initialization(fv1, fv2, prev, window, ratio, ori)
    assert r == rr and (m12 == rm12).all() and mine.tobytes() == rprev.tobytes() and r > 5
    r2, m12b = m.SearchForInitialization(_frame(k1, d1, sf), _frame(k2, d2, sf), mine, window)
    rr2, rm12b, _ = rp.search_for_initialization(fv1, fv2, rprev, window, ratio, ori)
    assert r2 == rr2 and (m12b == rm12b).all()


@pytest.mark.parametrize("seed,n,th,orb_dist,ori", [(1, 4000, 10.0, 100, True), (2, 2500, 15.0, 64, True), (3, 3000, 6.0, 90, False)])
def test_reloc_projection_vs_reference(tv, seed, n, th, orb_dist, ori):
    from orbslam3_cpp_fork_b200 import ORBmatcher
    k, d, sf = tv["k1"], tv["d1"], tv["sf"]
    q = point_queries(k, d, sf, n, seed, jitter=6.0)
    found = (np.random.default_rng(seed).random(n) < 0.1).astype(np.uint8)          # sAlreadyFound
    F = _frame(k, d, sf)
    F.mvpMapPoints[q["bound"] > 0] = 10 ** 6                      # bound on entry
    m = ORBmatcher(0.9, ori)
    r, best, rot, keepb = m.SearchByProjectionReloc(F, q["desc"], q["u"], q["v"], q["level"], q["angle"],
                                                    q["active"] & (1 - found), th, orb_dist)
    fv, keep = op.make_frame_view(k, d, None, q["bound"], W, H, sf)
    rr, rbind = rp.search_by_projection_reloc(fv, q["desc"], q["u"], q["v"], q["level"], q["angle"], q["active"], found, th,
                                              orb_dist, ori)
    kept = (best >= 0) & (np.isin(rot, keepb) if ori else True)
    assert r == rr and r > 300
    assert (rbind == binding_from_best(best, kept, q["bound"])).all()


@pytest.mark.parametrize("seed,n,th,ratio,with_kfs", [(4, 4000, 10, 1.0, False), (5, 3000, 5, 1.5, True), (6, 2000, 3, 0.8, False)])
def test_sim3_projection_vs_reference(tv, seed, n, th, ratio, with_kfs):
    from orbslam3_cpp_fork_b200 import ORBmatcher
    k, d, sf = tv["k2"], tv["d2"], tv["sf"]
    q = point_queries(k, d, sf, n, seed)
    m = ORBmatcher()
    r, best = m.SearchByProjectionSim3(_frame(k, d, sf), q["bound"], q["desc"], q["u"], q["v"], q["level"],
                                       q["active"] & in_image(q["u"], q["v"]), th, ratio)
    fv, keep = op.make_frame_view(k, d, None, q["bound"], W, H, sf)
    rr, rmatched = rp.search_by_projection_sim3(fv, q["desc"], q["u"], q["v"], q["level"], q["active"], th, ratio, with_kfs)
    assert r == rr and r > 100
    assert (rmatched == binding_from_best(best, best >= 0, q["bound"])).all()


@pytest.mark.parametrize("seed,n,th,sim3", [(7, 4000, 3.0, 0), (8, 3000, 4.0, 1), (9, 3000, 2.5, 0)])
def test_fuse_vs_reference(tv, seed, n, th, sim3):
    from orbslam3_cpp_fork_b200 import ORBmatcher
    k, d, sf = tv["k1"], tv["d1"], tv["sf"]
    q = point_queries(k, d, sf, n, seed, jitter=3.0)
    z, ur = fuse_depths(q, 40.0, seed)
    u_right = q["u_right"] if seed != 9 else None                 # seed 9: monocular KeyFrame (5.99 gate only)
    active = q["active"] & in_image(q["u"], q["v"])
    check = not sim3
    m = ORBmatcher()
    r, best = m.Fuse(_frame(k, d, sf, u_right), tv["inv_sigma2"] if check else None, q["desc"], q["u"], q["v"],
                     ur if check else None, q["level"], active, th, check)
    fv, keep = op.make_frame_view(k, d, u_right, None, W, H, sf, mb=0.1, mbf=40.0)
    rr, rbest, _ = rp.fuse(fv, q["desc"], q["u"], q["v"], z, q["level"], q["active"], th, sim3)
    assert r == rr and (best == rbest).all() and r > 100


@pytest.mark.parametrize("seed,th", [(10, 7.5), (11, 3.0)])
def test_search_by_sim3_vs_reference(tv, seed, th):
    from orbslam3_cpp_fork_b200 import ORBmatcher
    a, b = sim3_case(tv, seed)
    sf = tv["sf"]
    a_in = list(a); b_in = list(b)
    a_in[4] = a[4] & in_image(a[1], a[2]); b_in[4] = b[4] & in_image(b[1], b[2])
    m = ORBmatcher()
    r, m12 = m.SearchBySim3(_frame(tv["k1"], tv["d1"], sf), _frame(tv["k2"], tv["d2"], sf), *a_in, *b_in, th)
    fv1, k1 = op.make_frame_view(tv["k1"], tv["d1"], None, None, W, H, sf)
    fv2, k2 = op.make_frame_view(tv["k2"], tv["d2"], None, None, W, H, sf)
    rr, rm12 = rp.search_by_sim3(fv1, fv2, *a, *b, th)
    assert r == rr and (m12 == rm12).all() and r > 100


@pytest.mark.parametrize("pose_kind", ["sideways", "epipole_in_image"])
@pytest.mark.parametrize("seed,only_stereo,coarse,ori", [(16, False, False, True), (17, True, False, True), (18, False, True, True)])
def test_search_for_triangulation_vs_reference(tv, pose_kind, seed, only_stereo, coarse, ori):
    from orbslam3_cpp_fork_b200 import ORBmatcher
    c = triangulation_case(tv, seed)
    sf = tv["sf"]
    fv1, a = op.make_frame_view(tv["k1"], tv["d1"], c["ur1"], c["has1"], W, H, sf)
    fv2, b = op.make_frame_view(tv["k2"], tv["d2"], c["ur2"], c["has2"], W, H, sf)
    rr, rm12, ep, F12 = rp.search_for_triangulation(fv1, c["has1"], c["fv1"], fv2, c["has2"], c["fv2"],
                                                    triangulation_poses(pose_kind), only_stereo, coarse, ori)
    m = ORBmatcher(0.6, ori)
    r, m12 = m.SearchForTriangulation(_frame(tv["k1"], tv["d1"], sf, c["ur1"]), _frame(tv["k2"], tv["d2"], sf, c["ur2"]),
                                      c["has1"], c["has2"], c["fv1"], c["fv2"], F12.reshape(3, 3), ep, tv["sigma2"], only_stereo, coarse)
    assert r == rr and (m12 == rm12).all()
    if pose_kind == "sideways" and not only_stereo:
        assert r > 15
