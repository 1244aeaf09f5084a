"""CPU: the oracle's restatement of the ORBmatcher searches outside the tracking path (SURVEY.md §8f row N3:
oracle/search_oracle.cpp, oracle/bow_oracle.cpp) against the reference's OWN functions — the whole unmodified
src/ORBmatcher.cc compiled into oracle/_ref, driven through oracle/ref_compat/ref_capi_search.cpp.  Bit for bit on
every match index and count.

The reference functions take KeyFrame / MapPoint objects and run a geometric prologue (pose transform, projection,
distance range, viewing angle, PredictScale); the oracle and libcuorb take the prologue's results (u, v, level,
active).  tests/ref_search_cases.py holds how one is turned into the other; the harness places every MapPoint so
that the reference's own prologue reproduces exactly those numbers."""
import numpy as np
import pytest

from oracle import oracle_py as op
from oracle import ref_py as rp
from orbslam3_cpp_fork_b200.matcher import join_feature_vectors
from tests.bow_cases import make_pair
from tests.ref_search_cases import (binding_from_best, bow_views, fuse_bookkeeping, fuse_depths, has_mp_codes, in_image,
                                    triangulation_poses)
from tests.search_cases import H, W, point_queries, sim3_case, triangulation_case, two_views

pytestmark = pytest.mark.skipif(not rp.available(), reason="oracle/_ref not built and no reference checkout")


@pytest.fixture(scope="module")
def tv():
    return two_views(5, nf=800)


@pytest.mark.parametrize("seed,tie_heavy,mode,ratio,ori", [(1, False, 0, 0.75, True), (2, True, 1, 0.75, True),
                                                           (5, True, 0, 0.6, False), (7, False, 1, 0.9, True)])
def test_search_by_bow_equals_reference(seed, tie_heavy, mode, ratio, ori):
    c = make_pair(seed, n1=700, n2=640, n_nodes=60, tie_heavy=tie_heavy)
    fv1, fv2, keep = bow_views(c)
    h1 = has_mp_codes(c["valid1"], seed)                          # 0 none, 1 good MapPoint, 2 bad MapPoint
    h2 = None if mode == 0 else has_mp_codes(c["elig2"], seed + 1)
    elig2 = np.ones(len(c["desc2"]), np.uint8) if mode == 0 else (h2 == 1).astype(np.uint8)
    off1, idx1, off2, idx2 = join_feature_vectors(c["fv1"], c["fv2"])
    n, m = op.search_by_bow(c["desc1"], h1 == 1, c["angle1"], c["desc2"], elig2, c["angle2"], off1, idx1, off2, idx2,
                            mode, ratio, ori)
    rn, rm = rp.search_by_bow(mode, fv1, h1, c["fv1"], fv2, h2, c["fv2"], ratio, ori)
    assert rn == n and (rm == m).all() and n > 10


@pytest.mark.parametrize("tie_heavy,window,ratio,ori", [(False, 60, 0.9, True), (True, 60, 0.9, False), (True, 100, 0.9, True),
                                                       (False, 100, 0.8, True)])
def test_search_for_initialization_equals_reference(tie_heavy, window, ratio, ori):
    t = two_views(6, nf=500, tie_heavy=tie_heavy)
    fv1, a = op.make_frame_view(t["k1"], t["d1"], None, None, W, H, t["sf"])
    fv2, b = op.make_frame_view(t["k2"], t["d2"], None, None, W, H, t["sf"])
    prev = np.stack([t["k1"]["x"], t["k1"]["y"]], 1).astype(np.float32)
    n, m, pm = op.search_for_initialization(fv1, fv2, prev, window, ratio, ori)
    rn, rm, rpm = rp.search_for_initialization(fv1, fv2, prev, window, ratio, ori)
    assert rn == n and (rm == m).all() and rpm.tobytes() == pm.tobytes() and n > 3
    # second round on the updated vbPrevMatched, as Tracking::MonocularInitialization feeds it back
    n2, m2, pm2 = op.search_for_initialization(fv1, fv2, pm, window, ratio, ori)
    rn2, rm2, rpm2 = rp.search_for_initialization(fv1, fv2, rpm, window, ratio, ori)
    assert rn2 == n2 and (rm2 == m2).all() and rpm2.tobytes() == pm2.tobytes()


@pytest.mark.parametrize("seed,th,orb_dist,ori", [(3, 10.0, 80, True), (4, 3.0, 64, True), (5, 10.0, 100, False)])
def test_search_by_projection_reloc_equals_reference(tv, seed, th, orb_dist, ori):
    k, d, sf = tv["k1"], tv["d1"], tv["sf"]
    q = point_queries(k, d, sf, 600, seed)
    found = (np.random.default_rng(seed).random(600) < 0.1).astype(np.uint8)        # sAlreadyFound
    fv, keep = op.make_frame_view(k, d, None, q["bound"], W, H, sf)
    n, best, rot, keepb = op.search_by_projection_reloc(fv, q["desc"], q["u"], q["v"], q["level"], q["angle"],
                                                        q["active"] & (1 - found), th, orb_dist, ori)
    fv, keep = op.make_frame_view(k, d, None, q["bound"], W, H, sf)                  # the oracle marks bound_with_obs
    rn, rbind = rp.search_by_projection_reloc(fv, q["desc"], q["u"], q["v"], q["level"], q["angle"], q["active"], found,
                                              th, orb_dist, ori)
    kept = (best >= 0) & (np.isin(rot, keepb) if ori else True)
    assert rn == n and n > 60
    assert (rbind == binding_from_best(best, kept, q["bound"])).all()


@pytest.mark.parametrize("seed,th,ratio,with_kfs", [(3, 10, 1.0, False), (4, 6, 1.5, False), (5, 10, 1.0, True), (6, 4, 2.0, True)])
def test_search_by_projection_sim3_equals_reference(tv, seed, th, ratio, with_kfs):
    k, d, sf = tv["k1"], tv["d1"], tv["sf"]
    q = point_queries(k, d, sf, 600, seed)
    active = q["active"] & in_image(q["u"], q["v"])                                   # KeyFrame::IsInImage (:468)
    fv, keep = op.make_frame_view(k, d, None, q["bound"], W, H, sf)
    n, best = op.search_by_projection_sim3(fv, q["desc"], q["u"], q["v"], q["level"], active, th, ratio)
    fv, keep = op.make_frame_view(k, d, None, q["bound"], W, H, sf)
    rn, rmatched = rp.search_by_projection_sim3(fv, q["desc"], q["u"], q["v"], q["level"], q["active"], th, ratio, with_kfs)
    assert rn == n and n > 40
    assert (rmatched == binding_from_best(best, best >= 0, q["bound"])).all()


@pytest.mark.parametrize("seed,th,sim3", [(3, 3.0, 0), (4, 4.0, 0), (3, 3.0, 1), (5, 2.5, 1)])
def test_fuse_equals_reference(tv, seed, th, sim3):
    k, d, sf = tv["k1"], tv["d1"], tv["sf"]
    q = point_queries(k, d, sf, 600, seed)
    mbf = 40.0
    z, ur = fuse_depths(q, mbf, seed)                              # ur = u - bf * (1 / z), the reference's expression (:1208)
    active = q["active"] & in_image(q["u"], q["v"])
    fv, keep = op.make_frame_view(k, d, q["u_right"], q["bound"], W, H, sf, mb=0.1, mbf=mbf)
    n, best = op.fuse(fv, tv["inv_sigma2"], q["desc"], q["u"], q["v"], ur, q["level"], active, th, not sim3)
    rng = np.random.default_rng(seed)
    nobs_kf = rng.integers(1, 6, len(k)).astype(np.int32); nobs_q = rng.integers(1, 6, 600).astype(np.int32)
    rn, rbest, outcome = rp.fuse(fv, q["desc"], q["u"], q["v"], z, q["level"], q["active"], th, sim3, nobs_kf, nobs_q)
    assert (rbest == best).all() and rn == n == (best >= 0).sum() and n > 30
    # what the reference then does with the chosen feature (the shim leaves this bookkeeping to its caller and
    # INTEGRATION.md states the rule), replayed from best[] alone
    want = fuse_bookkeeping(best, q["bound"], nobs_kf, nobs_q, sim3)
    assert (outcome == want).all() and len(set(want.tolist())) == (3 if sim3 else 4)


@pytest.mark.parametrize("seed,th", [(11, 7.5), (12, 4.0)])
def test_search_by_sim3_equals_reference(tv, seed, th):
    a, b = sim3_case(tv, seed)
    fv1, k1 = op.make_frame_view(tv["k1"], tv["d1"], None, None, W, H, tv["sf"])
    fv2, k2 = op.make_frame_view(tv["k2"], tv["d2"], None, None, W, H, tv["sf"])
    a_in = list(a); b_in = list(b)
    a_in[4] = a[4] & in_image(a[1], a[2]); b_in[4] = b[4] & in_image(b[1], b[2])    # IsInImage (:1524 / :1608)
    n, m12 = op.search_by_sim3(fv1, fv2, *a_in, *b_in, th)
    rn, rm12 = rp.search_by_sim3(fv1, fv2, *a, *b, th)
    assert rn == n and (rm12 == m12).all() and n > 30


@pytest.mark.parametrize("pose_kind", ["sideways", "epipole_in_image"])
@pytest.mark.parametrize("only_stereo,coarse,ori", [(False, False, True), (True, False, False), (False, True, True)])
def test_search_for_triangulation_equals_reference(tv, pose_kind, only_stereo, coarse, ori):
    c = triangulation_case(tv, 4)
    pose1 = triangulation_poses(pose_kind)
    fv1, a = op.make_frame_view(tv["k1"], tv["d1"], c["ur1"], c["has1"], W, H, tv["sf"])
    fv2, b = op.make_frame_view(tv["k2"], tv["d2"], c["ur2"], c["has2"], W, H, tv["sf"])
    rn, rm12, ep, F12 = rp.search_for_triangulation(fv1, c["has1"], c["fv1"], fv2, c["has2"], c["fv2"], pose1,
                                                    only_stereo, coarse, ori)
    off1, idx1, off2, idx2 = join_feature_vectors(c["fv1"], c["fv2"])
    n, m12 = op.search_for_triangulation(fv1, fv2, off1, idx1, off2, idx2, F12, ep, tv["sigma2"], only_stereo, coarse, ori)
    assert rn == n and (rm12 == m12).all()
    if pose_kind == "sideways" and not only_stereo:
        assert n > 8
    if pose_kind == "epipole_in_image":
        assert 0 < ep[0] < W and 0 < ep[1] < H


@pytest.mark.parametrize("seed,tie_heavy,ratio,ori", [(1, False, 0.75, True), (2, True, 0.75, True), (5, True, 0.6, False),
                                                      (7, False, 0.9, True)])
def test_search_by_bow_rig_equals_reference(seed, tie_heavy, ratio, ori):
    """SearchByBoW(KeyFrame, Frame) with F.Nleft != -1 (src/ORBmatcher.cc:306-331, :371-396): a best / second-best pair
    per camera, the right match nested in the left one's threshold test with its ratio test disabled."""
    from tests.ref_search_cases import bow_rig_views
    c = make_pair(seed, n1=700, n2=640, n_nodes=60, tie_heavy=tie_heavy)
    nleft1, nleft2 = 420, 350
    (kl, kr, fl, fr), keep = bow_rig_views(c, nleft1, nleft2)
    h1 = has_mp_codes(c["valid1"], seed)
    off1, idx1, off2, idx2 = join_feature_vectors(c["fv1"], c["fv2"])
    n, ml, mr = op.search_by_bow_rig(c["desc1"], h1 == 1, c["angle1"], c["desc2"], nleft2, c["angle2"], off1, idx1, off2, idx2,
                                     ratio, ori)
    rn, rml, rmr = rp.search_by_bow_rig(kl, kr, h1, c["fv1"], fl, fr, c["fv2"], ratio, ori)
    assert rn == n and (rml == ml).all() and (rmr == mr).all()
    assert (ml >= 0).sum() > 10 and (mr >= 0).sum() > 10 and ((ml >= 0) & (mr >= 0)).sum() > 0
    assert (ml[ml >= 0] < nleft2).all() and (mr[mr >= 0] >= nleft2).all()


@pytest.mark.parametrize("right,seed,th", [(0, 3, 3.0), (1, 4, 3.0), (1, 5, 4.0)])
def test_fuse_on_rig_keyframe_equals_reference(right, seed, th):
    """Fuse(pKF, vpMapPoints, th, bRight) on a two-camera rig KeyFrame (src/ORBmatcher.cc:1156-1164, :1236, :1270-1272) is
    the flat search on the chosen camera's features (no mvuRight: only the 5.99 gate) plus NLeft on the index."""
    from tests.rig_cases import W as RW, H as RH, frame_views, rig_views
    rv = rig_views(7)
    sf = rv["sf"]
    inv_sigma2 = (np.float32(1.0) / (sf * sf)).astype(np.float32)
    k, d = (rv["k2"], rv["d2"]) if right else (rv["k1"], rv["d1"])
    q = point_queries(k, d, sf, 600, seed)
    z, ur = fuse_depths(q, 40.0, seed)
    fv, keep = op.make_frame_view(k, d, None, None, RW, RH, sf, mb=0.1, mbf=40.0)
    n, best = op.fuse(fv, inv_sigma2, q["desc"], q["u"], q["v"], ur, q["level"], q["active"] & in_image(q["u"], q["v"]), th, True)
    fvl, fvr, kk = frame_views(rv)
    rn, rbest = rp.fuse_rig(fvl, fvr, right, q["desc"], q["u"], q["v"], z, q["level"], q["active"], th)
    off = len(rv["k1"]) if right else 0
    assert rn == n and (rbest == np.where(best >= 0, best + off, -1)).all() and n > 100
