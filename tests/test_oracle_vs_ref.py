"""CPU: the restated oracle (oracle/_build/liborb_oracle.so) against the reference's OWN code
(oracle/_ref/liborb_ref.so = unmodified src/ORBextractor.cc + src/ORBmatcher.cc + the bodies lifted
verbatim out of src/Frame.cc, compiled over oracle/ref_compat).  Bit for bit: this is what pins the
oracle above the OpenCV boundary — std::list / std::sort quadtree, cell loop, output scatter, stereo
association, grid, projection searches."""
import glob
import os

import numpy as np
import pytest

from oracle import oracle_py as op
from oracle import ref_py as rp
from orbslam3_cpp_fork_b200 import synth
from tests.matching_cases import extracted_frame, frustum_case, lastframe_case, mappoint_case
from tests.test_oracle_golden import GOLDEN, golden_image

pytestmark = pytest.mark.skipif(not rp.available(), reason="oracle/_ref not built and no reference checkout")

CASES = [(752, 480, 1000, 1, (0, 0), None), (752, 480, 1000, 2, (0, 1000), None), (640, 480, 1000, 3, (0, 0), None),
         (1241, 376, 2000, 4, (0, 0), None), (320, 240, 500, 5, (100, 200), None), (752, 480, 5000, 8, (0, 0), None),
         (752, 480, 1000, 6, (0, 0), "black"), (752, 480, 1000, 6, (0, 0), "flat"), (752, 480, 1000, 6, (0, 0), "noise")]


def assert_same(a, b):
    mo, ko, do = a
    mr, kr, dr = b
    assert mo == mr and len(ko) == len(kr)
    assert ko.tobytes() == kr.tobytes()          # every field of every cv::KeyPoint, bit for bit
    assert (do == dr).all()


def test_ctor_tables():
    for args in ((1000, 1.2, 8, 20, 7), (2000, 1.2, 8, 20, 7), (5000, 1.2, 8, 20, 7), (800, 1.5, 5, 20, 7),
                 (1000, 1.1, 12, 20, 7), (300, 1.3, 6, 12, 12)):
        o, r = op.OracleExtractor(*args), rp.RefExtractor(*args)
        for a, b in zip(o.tables(), r.tables()):
            assert a.tobytes() == b.tobytes()
        assert (o.umax() == r.umax()).all()


@pytest.mark.parametrize("w,h,nf,seed,lap,kind", CASES)
def test_extractor_equals_reference(w, h, nf, seed, lap, kind):
    img = synth.degenerate_image(w, h, kind, seed) if kind else synth.synth_image(w, h, seed)
    o, r = op.OracleExtractor(nf), rp.RefExtractor(nf)
    assert_same(o.extract(img, lap), r.extract(img, lap))
    for lv in range(8):
        assert (o.plane(lv, 0) == r.plane(lv, 0)).all(), f"bordered pyramid level {lv}"
        assert o.level_keys(lv).tobytes() == r.level_keys(lv).tobytes(), f"allKeypoints[{lv}]"


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_reference_reproduces_cv2_goldens(path):
    """The goldens were made with the real cv2 primitives (tests/golden/make_golden.py); the reference's
    own operator() over the restated primitives lands on the same keypoints and descriptors."""
    g = np.load(path)
    img = golden_image(g)
    r = rp.RefExtractor(int(g["nfeatures"]))
    mono, k, d = r.extract(img, tuple(int(v) for v in g["lap"]))
    assert mono == int(g["mono"]) and len(k) == len(g["kps"])
    for f in ("x", "y", "size", "angle", "response", "octave", "class_id"):
        assert (k[f] == g["kps"][f]).all(), f
    assert (d == g["desc"]).all()


@pytest.mark.parametrize("w,h,nf,scale,nlevels,ini,mn,seed", [
    (641, 479, 1000, 1.2, 8, 20, 7, 91), (1280, 720, 1500, 1.2, 8, 20, 7, 92), (752, 480, 1000, 1.1, 12, 20, 7, 94),
    (752, 480, 800, 1.5, 5, 20, 7, 95), (400, 300, 300, 1.3, 6, 12, 12, 96)])
def test_other_geometries(w, h, nf, scale, nlevels, ini, mn, seed):
    img = synth.synth_image(w, h, seed)
    o, r = op.OracleExtractor(nf, scale, nlevels, ini, mn), rp.RefExtractor(nf, scale, nlevels, ini, mn)
    assert_same(o.extract(img), r.extract(img))


def test_empty_image_returns_minus_one():
    assert rp.RefExtractor(500).extract(np.zeros((0, 0), np.uint8))[0] == -1   # src/ORBextractor.cc:1060


def test_distribute_octtree_ties_and_overshoot():
    """DistributeOctTree alone (src/ORBextractor.cc:556-753) on tie-heavy candidate sets: equal responses,
    clustered points, more / fewer candidates than N, several root nodes."""
    o, r = op.OracleExtractor(1000), rp.RefExtractor(1000)
    rng = np.random.default_rng(17)
    for case in range(60):
        H = int(rng.integers(60, 500)); W = int(rng.integers(max(80, H), 1300))   # nIni = round(W/H) >= 1 (0 is UB, :566)
        n = int(rng.integers(1, 6000)); N = int(rng.integers(1, 600))
        keys = np.zeros(n, op.KP_DTYPE)
        if case % 3 == 0:      # clustered
            cx, cy = rng.uniform(0, W, 8), rng.uniform(0, H, 8)
            sel = rng.integers(0, 8, n)
            keys["x"] = np.clip(np.floor(cx[sel] + rng.normal(0, 12, n)), 0, W - 1)
            keys["y"] = np.clip(np.floor(cy[sel] + rng.normal(0, 12, n)), 0, H - 1)
        else:
            keys["x"] = rng.integers(0, W, n); keys["y"] = rng.integers(0, H, n)
        keys["response"] = rng.integers(7, 12 if case % 2 else 255, n)   # few distinct values -> many ties
        keys["size"] = 7; keys["angle"] = -1; keys["class_id"] = -1
        a = o.octree(keys, 0, W, 0, H, N); b = r.octree(keys, 0, W, 0, H, N)
        assert a.tobytes() == b.tobytes(), f"case {case}: n={n} N={N} {W}x{H}"


@pytest.mark.parametrize("w,h,nf,seed,fx,mbf", [(752, 480, 1000, 11, 458.654, 47.906), (1241, 376, 2000, 12, 718.856, 386.1448),
                                               (320, 240, 500, 7, 435.2, 47.9)])
def test_stereo_equals_reference(w, h, nf, seed, fx, mbf):
    """Frame::ComputeStereoMatches: mvuRight / mvDepth bit-identical floats."""
    L, R = synth.synth_stereo_pair(w, h, seed)
    oL, oR, rL, rR = op.OracleExtractor(nf), op.OracleExtractor(nf), rp.RefExtractor(nf), rp.RefExtractor(nf)
    _, k0, d0 = oL.extract(L); _, k1, d1 = oR.extract(R)
    rL.extract(L); rR.extract(R)
    mb = mbf / fx
    so = op.stereo(oL, oR, k0, d0, k1, d1, mb, mbf)
    u, d = rp.stereo(rL, rR, k0, d0, k1, d1, mb, mbf)
    assert (so["u_right"] >= 0).sum() > 50
    assert so["u_right"].tobytes() == u.tobytes() and so["depth"].tobytes() == d.tobytes()


def test_hamming_equals_reference():
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, (200, 32), dtype=np.uint8); b = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    for i in range(200):
        assert op.hamming(a[i], b[i]) == rp.hamming(a[i], b[i]) == int(np.unpackbits(a[i] ^ b[i]).sum())


def test_grid_equals_reference():
    """AssignFeaturesToGrid / GetFeaturesInArea (src/Frame.cc:524-553, 788-877) on an extracted frame,
    including queries that hang over the image border and level-filtered queries."""
    e, k, d = extracted_frame()
    sf = e.tables()[0]
    fv, keep = op.make_frame_view(k, d, None, None, 752, 480, sf)
    F = op.lib()
    rng = np.random.default_rng(9)
    import ctypes as C
    for _ in range(300):
        x, y = float(np.float32(rng.uniform(-20, 772))), float(np.float32(rng.uniform(-20, 500)))
        r = float(np.float32(rng.uniform(1, 60)))
        lo, hi = int(rng.integers(-1, 7)), int(rng.integers(-1, 8))
        ref = rp.features_in_area(fv, x, y, r, lo, hi)
        out = np.zeros(4096, np.int32)
        n = F.oo_features_in_area(C.byref(fv), x, y, r, lo, hi, op._p(out), 4096)
        assert n == len(ref) and (out[:n] == ref).all()      # same members in the same visiting order
        # independent statement of the contract (src/Frame.cc:830-857)
        chk = (lo > 0) or (hi >= 0)
        ok = (np.abs(k["x"] - np.float32(x)) < np.float32(r)) & (np.abs(k["y"] - np.float32(y)) < np.float32(r))
        if chk:
            ok &= k["octave"] >= lo
            if hi >= 0:
                ok &= k["octave"] <= hi
        assert set(ref.tolist()) == set(np.nonzero(ok)[0].tolist())


def replay_mappoints(best, bound, n_feat):
    final = np.where(np.asarray(bound) > 0, -2, -1).astype(np.int32) if bound is not None else np.full(n_feat, -1, np.int32)
    for i, b in enumerate(best):
        if b >= 0:
            final[b] = i
    return final


@pytest.mark.parametrize("M,th,with_uright,far", [(50000, 3.0, False, False), (50000, 1.0, True, True), (3000, 3.0, True, False)])
def test_search_by_projection_mappoints_equals_reference(M, th, with_uright, far):
    e, k, d = extracted_frame()
    sf = e.tables()[0]
    c = mappoint_case(k, d, sf, M, 77 + M, with_uright=with_uright)
    fv, keep = op.make_frame_view(k, d, c["u_right"], c["bound"], 752, 480, sf, mb=0.1, mbf=40.0)
    args = (c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"], c["depth"], c["flags"], th, far, 60.0, 0.8)
    n_o, best_o = op.search_by_projection(fv, *args)
    n_r, best_r, final_r = rp.search_by_projection(fv, *args)
    assert n_r > 100
    assert n_o == n_r and (best_o == best_r).all()
    assert (replay_mappoints(best_o, c["bound"], len(k)) == final_r).all()
    # a case where order matters was really exercised: some feature is claimed by more than one MapPoint
    claimed = best_r[best_r >= 0]
    assert len(np.unique(claimed)) < len(claimed) or M < 10000


@pytest.mark.parametrize("fwd,bwd,th", [(False, False, 7.0), (True, False, 15.0), (False, True, 7.0)])
def test_search_by_projection_lastframe_equals_reference(fwd, bwd, th):
    e, k, d = extracted_frame(seed=33)
    sf = e.tables()[0]
    c = lastframe_case(k, d, 5)
    z = (np.float32(1.0) / c["inv_z"]).astype(np.float32)
    inv_z = (1.0 / z.astype(np.float64)).astype(np.float32)        # what `1.0 / x3Dc(2)` gives (:1715)
    fv, keep = op.make_frame_view(k, d, c["u_right"], None, 752, 480, sf, mb=0.1, mbf=40.0)
    n_o, best_o, rot_o, keep_o = op.search_by_projection_frame(fv, c["desc"], c["u"], c["v"], inv_z, c["octave"],
                                                              c["angle"], c["flags"], th, fwd, bwd, True)
    n_r, final_r = rp.search_by_projection_frame(fv, c["desc"], c["u"], c["v"], z, c["octave"], c["angle"],
                                                 c["flags"], th, fwd, bwd, True)
    # replay as the reference does: bind every accepted match in order (:1786), then NULL the features of
    # every non-kept histogram bin (:1889-1905)
    final = np.full(len(k), -1, np.int32)
    for i, b in enumerate(best_o):
        if b >= 0:
            final[b] = i
    for i, b in enumerate(best_o):
        if b >= 0 and rot_o[i] not in keep_o:
            final[b] = -1
    assert n_r > 50
    assert n_o == n_r
    assert (final == final_r).all()


@pytest.mark.parametrize("seed,limit", [(1, 0.5), (2, 0.5), (3, 0.8)])
def test_is_in_frustum_equals_reference(seed, limit):
    """Frame::isInFrustum (src/Frame.cc:644-707) + MapPoint::PredictScale + Pinhole::project: the restatement
    against the reference's own lifted bodies, every output bit for bit."""
    cam, P, N, min_d, max_d = frustum_case(20000, seed)
    a = op.is_in_frustum(cam, P, N, min_d, max_d, limit)
    b = rp.is_in_frustum(cam, P, N, min_d, max_d, limit)
    assert 2000 < a["count"] < 18000
    assert a["count"] == b["count"]
    for k in ("in_view", "proj_x", "proj_y", "proj_xr", "level", "view_cos", "depth"):
        assert a[k].tobytes() == b[k].tobytes(), k
    assert len(np.unique(a["level"][a["in_view"] > 0])) == 8          # every scale level is predicted somewhere
