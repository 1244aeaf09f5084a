"""GPU (-m gpu): the C++ drop-in shim (shim/ORBextractor.h, shim/ORBmatcher.h) compiled with g++
against libcuorb.so, driven like the reference's stereo Frame constructor (src/Frame.cc:202-225),
and compared bit for bit with the CPU oracle."""
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle_py as op
from orbslam3_cpp_fork_b200 import _lib, build, synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_shim_stereo_frame(tmp_path):
    build.build_libcuorb()
    exe = str(tmp_path / "shim_check")
    pkg = os.path.join(ROOT, "orbslam3_cpp_fork_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", os.path.join(ROOT, "tests", "native", "shim_check.cpp"), "-o", exe,
                           "-L" + pkg, "-l:libcuorb.so", "-Wl,-rpath," + pkg])
    w, h, nf = 640, 480, 1000
    fx, mbf = 525.0, 40.0
    mb = mbf / fx
    left, right = synth.synth_stereo_pair(w, h, seed=21)
    left.tofile(tmp_path / "l.raw"); right.tofile(tmp_path / "r.raw")
    out = str(tmp_path / "out.bin")
    subprocess.check_call([exe, str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(w), str(h), str(nf), repr(mb), repr(mbf), out])
    raw = open(out, "rb").read()
    hdr = np.frombuffer(raw, np.int32, 8)
    nl, nr, mono_l, mono_r, selfdist, d01, pw, ph = [int(v) for v in hdr]
    off = 32
    kl = np.frombuffer(raw, _lib.KP_DTYPE, nl, off); off += 28 * nl
    dl = np.frombuffer(raw, np.uint8, 32 * nl, off).reshape(nl, 32); off += 32 * nl
    kr = np.frombuffer(raw, _lib.KP_DTYPE, nr, off); off += 28 * nr
    dr = np.frombuffer(raw, np.uint8, 32 * nr, off).reshape(nr, 32); off += 32 * nr
    ur = np.frombuffer(raw, np.float32, nl, off); off += 4 * nl
    dp = np.frombuffer(raw, np.float32, nl, off); off += 4 * nl
    pyr1 = np.frombuffer(raw, np.uint8, pw * ph, off).reshape(ph, pw)

    oL, oR = op.OracleExtractor(nf, 1.2, 8, 20, 7), op.OracleExtractor(nf, 1.2, 8, 20, 7)
    m0, k0, d0 = oL.extract(left); m1, k1, d1 = oR.extract(right)
    assert (nl, nr, mono_l, mono_r) == (len(k0), len(k1), m0, m1)
    assert (kl == k0).all() and (dl == d0).all() and (kr == k1).all() and (dr == d1).all()
    so = op.stereo(oL, oR, k0, d0, k1, d1, np.float32(mb), np.float32(mbf))
    assert (ur == so["u_right"]).all() and (dp == so["depth"]).all()
    assert selfdist == 0 and d01 == int(np.unpackbits(d0[0] ^ d0[1]).sum())
    assert (pyr1 == oL.plane(1, 1)).all()


def test_cpp_shim_frame_prep(tmp_path):
    """shim/FramePrep.h: rectify + gray (+ fused extraction) of a BGR frame == oracle chain."""
    from tests.preproc_cases import color_frames, rectify_maps
    build.build_libcuorb()
    exe = str(tmp_path / "prep_check")
    pkg = os.path.join(ROOT, "orbslam3_cpp_fork_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", os.path.join(ROOT, "tests", "native", "prep_check.cpp"), "-o", exe,
                           "-L" + pkg, "-l:libcuorb.so", "-Wl,-rpath," + pkg])
    w, h, nf = 752, 480, 1000
    frame = color_frames(1, h, w, 3, 90)[0]
    mx, my = rectify_maps(w, h, 2)
    frame.tofile(tmp_path / "f.raw"); mx.tofile(tmp_path / "mx.raw"); my.tofile(tmp_path / "my.raw")
    out = str(tmp_path / "out.bin")
    subprocess.check_call([exe, str(tmp_path / "f.raw"), str(w), str(h), str(tmp_path / "mx.raw"), str(tmp_path / "my.raw"), str(nf), out])
    raw = open(out, "rb").read()
    n, mono = [int(v) for v in np.frombuffer(raw, np.int32, 2)]
    off = 8
    gray = np.frombuffer(raw, np.uint8, w * h, off).reshape(h, w); off += w * h
    k = np.frombuffer(raw, _lib.KP_DTYPE, n, off); off += 28 * n
    d = np.frombuffer(raw, np.uint8, 32 * n, off).reshape(n, 32)
    ref_gray = op.preprocess(frame, op.PIX_BGR, op.PRE_REMAP, mx, my)
    assert (gray == ref_gray).all()
    m0, k0, d0 = op.OracleExtractor(nf, 1.2, 8, 20, 7).extract(ref_gray)
    assert (n, mono) == (len(k0), m0)
    assert (k == k0).all() and (d == d0).all()


def test_cpp_shim_searches(tmp_path):
    """shim/ORBmatcher.h: SearchForInitialization (twice, as monocular initialisation does), the Sim3
    SearchByProjection, Fuse and SearchForTriangulation on reference-named structs == oracle."""
    from orbslam3_cpp_fork_b200.matcher import join_feature_vectors
    build.build_libcuorb()
    exe = str(tmp_path / "search_check")
    pkg = os.path.join(ROOT, "orbslam3_cpp_fork_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", os.path.join(ROOT, "tests", "native", "search_check.cpp"), "-o", exe,
                           "-L" + pkg, "-l:libcuorb.so", "-Wl,-rpath," + pkg])
    w, h, nf = 640, 480, 2000
    a, b = synth.synth_stereo_pair(w, h, seed=33)
    a.tofile(tmp_path / "a.raw"); b.tofile(tmp_path / "b.raw")
    out = str(tmp_path / "out.bin")
    subprocess.check_call([exe, str(tmp_path / "a.raw"), str(tmp_path / "b.raw"), str(w), str(h), str(nf), out])
    raw = open(out, "rb").read()
    na, nb, n_ini1, n_ini2, n_sim3, n_fuse, n_tri, _ = [int(v) for v in np.frombuffer(raw, np.int32, 8)]
    off = 32
    ka = np.frombuffer(raw, _lib.KP_DTYPE, na, off); off += 28 * na
    da = np.frombuffer(raw, np.uint8, 32 * na, off).reshape(na, 32); off += 32 * na
    kb = np.frombuffer(raw, _lib.KP_DTYPE, nb, off); off += 28 * nb
    db = np.frombuffer(raw, np.uint8, 32 * nb, off).reshape(nb, 32); off += 32 * nb
    ini1 = np.frombuffer(raw, np.int32, na, off); off += 4 * na
    ini2 = np.frombuffer(raw, np.int32, na, off); off += 4 * na
    prev = np.frombuffer(raw, np.float32, 2 * na, off).reshape(na, 2); off += 8 * na
    sim3 = np.frombuffer(raw, np.int32, na, off); off += 4 * na
    fused = np.frombuffer(raw, np.int32, na, off); off += 4 * na
    tri = np.frombuffer(raw, np.int32, na, off)

    o = op.OracleExtractor(nf, 1.2, 8, 20, 7)
    sf = np.asarray(o.tables()[0], np.float32)
    fva, keep_a = op.make_frame_view(ka, da, None, None, w, h, sf, mb=0.1, mbf=40.0)
    fvb, keep_b = op.make_frame_view(kb, db, None, None, w, h, sf, mb=0.1, mbf=40.0)
    p0 = np.stack([ka["x"], ka["y"]], 1).astype(np.float32)
    r1, m1, p1 = op.search_for_initialization(fva, fvb, p0, 100, 0.9, True)
    r2, m2, p2 = op.search_for_initialization(fva, fvb, p1, 100, 0.9, True)
    assert (n_ini1, n_ini2) == (r1, r2) and (ini1 == m1).all() and (ini2 == m2).all() and (prev == p2).all() and r1 > 20

    i = np.arange(na)
    level = np.minimum(ka["octave"] + (i & 1), 7).astype(np.int32)
    active = (i % 7 != 0).astype(np.uint8)
    u = (ka["x"] + np.float32(1.5)).astype(np.float32); v = (ka["y"] - np.float32(0.75)).astype(np.float32)
    bound = (i % 5 == 0).astype(np.uint8)
    fvs, keep_s = op.make_frame_view(ka, da, None, bound, w, h, sf, mb=0.1, mbf=40.0)
    rs, best = op.search_by_projection_sim3(fvs, da, u, v, level, active, 10, 1.0)
    ref = np.full(na, -1, np.int32)
    ref[best[best >= 0]] = i[best >= 0]
    assert n_sim3 == rs and (sim3 == ref).all() and rs > 100
    inv_sigma2 = (np.float32(1.0) / (sf * sf)).astype(np.float32)
    rf, bf = op.fuse(fva, inv_sigma2, da, u, v, (ka["x"] - np.float32(9.0)).astype(np.float32), level, active, 3.0, True)
    assert n_fuse == rf and (fused == bf).all() and rf > 100

    F12 = np.array([0, 1e-4, 0, -2e-4, 0, -1, 3e-4, 1, 0], np.float32)
    ep = np.array([w * 0.5, h * 0.5], np.float32)
    has_a = (i % 3 == 0).astype(np.uint8)
    fv1 = {}; fv2 = {}
    for j in range(na): fv1.setdefault(int(ka["y"][j]) // 10, []).append(j)
    for j in range(nb): fv2.setdefault(int(kb["y"][j]) // 10, []).append(j)
    off1, idx1, off2, idx2 = join_feature_vectors(fv1, fv2)
    fvt, keep_t = op.make_frame_view(ka, da, None, has_a, w, h, sf, mb=0.1, mbf=40.0)
    rt, mt = op.search_for_triangulation(fvt, fvb, off1, idx1, off2, idx2, F12, ep, (sf * sf).astype(np.float32), False, False, True)
    assert n_tri == rt and (tri == mt).all() and rt > 5


@pytest.mark.parametrize("motion,th,mono", [("none", 7.0, False), ("forward", 15.0, False), ("backward", 7.0, False),
                                            ("forward", 15.0, True)])
def test_cpp_shim_search_by_projection_lastframe(tmp_path, motion, th, mono):
    """shim/ORBmatcher.h SearchByProjection(Cur, Last, th, bMono) - Tracking::TrackWithMotionModel's call - with real
    poses and world points: the FINAL CurrentFrame.mvpMapPoints and the return value equal the reference's own
    function (oracle/_ref, src/ORBmatcher.cc:1684-1908), including features that are bound twice (a 0-observation
    MapPoint overwritten by a later one) and the NULLing of the non-kept rotation bins."""
    from oracle import ref_py as rp
    from tests.matching_cases import extracted_frame, lastframe_pose_case
    if not rp.available():
        pytest.skip("oracle/_ref/liborb_ref.so did not travel to this box")
    build.build_libcuorb()
    exe = str(tmp_path / "lastframe_check")
    pkg = os.path.join(ROOT, "orbslam3_cpp_fork_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-DCUORB_SHIM_USE_OPENCV",
                           "-I" + os.path.join(ROOT, "oracle", "ref_compat"),
                           os.path.join(ROOT, "tests", "native", "lastframe_check.cpp"), "-o", exe,
                           "-L" + pkg, "-l:libcuorb.so", "-Wl,-rpath," + pkg])
    e, k, d = extracted_frame(seed=33)
    sf = e.tables()[0]
    w, h = 752, 480
    c = lastframe_pose_case(k, d, 9, motion)
    mb, mbf = np.float32(0.1), np.float32(40.0)
    fv, keep = op.make_frame_view(k, d, c["u_right"], c["bound"], w, h, sf, mb=float(mb), mbf=float(mbf))
    n_r, final_r = rp.search_by_projection_frame_pose(fv, c["pose_cur"], c["pose_last"], c["K4"], c["desc"], c["world"],
                                                      c["octave"], c["angle"], c["has_mp"], c["outlier"], c["n_obs"], th, mono)
    case = str(tmp_path / "case.bin"); out = str(tmp_path / "out.bin")
    with open(case, "wb") as f:
        np.array([len(k), len(c["octave"]), len(sf), int(mono), 1], np.int32).tofile(f)
        np.concatenate([[th, mb, mbf, w, h], c["K4"], c["pose_cur"], c["pose_last"]]).astype(np.float32).tofile(f)
        np.ascontiguousarray(k).tofile(f); np.ascontiguousarray(d).tofile(f)
        c["u_right"].tofile(f); c["bound"].tofile(f); np.ascontiguousarray(sf, np.float32).tofile(f)
        c["world"].tofile(f); c["octave"].tofile(f); c["angle"].tofile(f); c["has_mp"].tofile(f); c["outlier"].tofile(f)
        c["n_obs"].tofile(f); c["desc"].tofile(f)
    subprocess.check_call([exe, case, out])
    res = np.fromfile(out, np.int32)
    n_s, final_s = int(res[0]), res[1:]
    assert n_r > 100
    assert n_s == n_r
    assert (final_s == final_r).all()
    # the order-dependent situation really occurred: fewer bound features than accepted matches + removed ones
    assert (final_r >= 0).sum() <= n_r


@pytest.mark.parametrize("which,seed", [("mappoints", 1), ("mappoints", 3), ("lastframe", 1), ("lastframe", 2)])
def test_cpp_shim_rig_branches(tmp_path, which, seed):
    """shim/ORBmatcher.h on two-camera rig Frames (Nleft != -1): SearchByProjection(F, vpMapPoints, ...) and
    SearchByProjection(Cur, Last, ...) leave the same F.mvpMapPoints (both cameras' slots, cross-bound stereo partners,
    rotation NULLing) and return the same count as the reference's own functions (src/ORBmatcher.cc:38-227, :1684-1908)."""
    from oracle import ref_py as rp
    from tests.rig_cases import H, W, frame_views, rig_lastframe, rig_mappoints, rig_views
    if not rp.available():
        pytest.skip("oracle/_ref/liborb_ref.so did not travel to this box")
    build.build_libcuorb()
    exe = str(tmp_path / "rig_check")
    pkg = os.path.join(ROOT, "orbslam3_cpp_fork_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-DCUORB_SHIM_USE_OPENCV",
                           "-I" + os.path.join(ROOT, "oracle", "ref_compat"),
                           os.path.join(ROOT, "tests", "native", "rig_check.cpp"), "-o", exe,
                           "-L" + pkg, "-l:libcuorb.so", "-Wl,-rpath," + pkg])
    rv = rig_views(9, nf=1200)
    fvl, fvr, keep = frame_views(rv)
    th, thfar, ratio, mb, mbf = (1.0 if seed == 1 else 3.0), 40.0, 0.8, 0.1, 40.0
    case = str(tmp_path / "case.bin"); out = str(tmp_path / "out.bin")
    if which == "mappoints":
        q = rig_mappoints(rv, 6000, seed, no_obs_frac=0.1 if seed == 1 else 0.5)
        n_r, final_r, _ = rp.search_by_projection_rig(fvl, fvr, rv["l2r"], rv["r2l"], q["desc"], q["px"], q["py"], q["level"],
                                                      q["view_cos"], q["pxr"], q["pyr"], q["level_r"], q["view_cos_r"], q["depth"],
                                                      q["flags"], th, seed == 3, thfar, ratio)
        hi = [len(rv["k1"]), len(rv["k2"]), len(rv["sf"]), 6000, int(seed == 3), 0, 1]
        hf = [th, thfar, ratio, mb, mbf, W, H, 0, 0]
        tail = [q["desc"], q["px"], q["py"], q["view_cos"], q["pxr"], q["pyr"], q["view_cos_r"], q["depth"], q["level"], q["level_r"],
                q["flags"]]
    else:
        q = rig_lastframe(rv, seed)
        fwd, th = (seed == 2), (15.0 if seed == 2 else 7.0)
        n_r, final_r = rp.search_by_projection_frame_rig(fvl, fvr, q["desc"], q["u"], q["v"], q["z"], q["trl"][0], q["trl"][1],
                                                         q["octave"], q["angle"], q["flags"], th, fwd, False, True)
        hi = [len(rv["k1"]), len(rv["k2"]), len(rv["sf"]), len(q["u"]), int(fwd), 0, 1]
        hf = [th, thfar, ratio, mb, mbf, W, H, q["trl"][0], q["trl"][1]]
        tail = [q["desc"], q["u"], q["v"], q["z"], q["angle"], q["octave"], q["flags"]]
    with open(case, "wb") as f:
        np.array(hi, np.int32).tofile(f); np.array(hf, np.float32).tofile(f)
        np.ascontiguousarray(rv["k1"]).tofile(f); np.ascontiguousarray(rv["k2"]).tofile(f)
        np.ascontiguousarray(rv["d1"]).tofile(f); np.ascontiguousarray(rv["d2"]).tofile(f)
        np.concatenate([rv["bound_l"], rv["bound_r"]]).astype(np.uint8).tofile(f)
        rv["l2r"].tofile(f); rv["r2l"].tofile(f); np.ascontiguousarray(rv["sf"], np.float32).tofile(f)
        for a in tail:
            np.ascontiguousarray(a).tofile(f)
    subprocess.check_call([exe, which, case, out])
    res = np.fromfile(out, np.int32)
    assert n_r > 300
    assert int(res[0]) == n_r
    assert (res[1:] == final_r).all()


@pytest.mark.parametrize("seed,mono_l,mono_r", [(1, 0, 0), (2, 300, 250), (3, 1195, 0)])
def test_cpp_shim_fisheye_stereo(tmp_path, seed, mono_l, mono_r):
    """shim/FrameOps.h ComputeStereoFishEyeMatches (lapping split, 2-NN + ratio test on the device, triangulation
    hand-off, stereo tables) against the reference's own function lifted into oracle/_ref (src/Frame.cc:1303-1355); both
    sides triangulate with the deterministic stand-in of oracle/ref_compat/fake_triangulate.h."""
    from oracle import ref_py as rp
    from tests.rig_cases import rig_views
    if not rp.available():
        pytest.skip("oracle/_ref/liborb_ref.so did not travel to this box")
    build.build_libcuorb()
    exe = str(tmp_path / "fisheye_check")
    pkg = os.path.join(ROOT, "orbslam3_cpp_fork_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-DCUORB_SHIM_USE_OPENCV",
                           "-I" + os.path.join(ROOT, "oracle", "ref_compat"),
                           os.path.join(ROOT, "tests", "native", "fisheye_check.cpp"), "-o", exe,
                           "-L" + pkg, "-l:libcuorb.so", "-Wl,-rpath," + pkg])
    rv = rig_views(9 + seed, nf=1200)
    k1, d1, k2, d2 = rv["k1"], rv["d1"], rv["k2"], rv["d2"]
    mono_l = min(mono_l, len(k1)); mono_r = min(mono_r, len(k2))
    sigma2 = (rv["sf"] * rv["sf"]).astype(np.float32)
    l2r, r2l, depth, p3d = rp.fisheye_stereo(k1, d1, mono_l, k2, d2, mono_r, sigma2)
    case = str(tmp_path / "case.bin"); out = str(tmp_path / "out.bin")
    with open(case, "wb") as f:
        np.array([len(k1), len(k2), mono_l, mono_r, len(sigma2)], np.int32).tofile(f)
        np.ascontiguousarray(k1).tofile(f); np.ascontiguousarray(k2).tofile(f)
        np.ascontiguousarray(d1).tofile(f); np.ascontiguousarray(d2).tofile(f); sigma2.tofile(f)
    subprocess.check_call([exe, case, out])
    raw = np.fromfile(out, np.uint8)
    nl, nr = len(k1), len(k2)
    s_l2r = raw[:4 * nl].view(np.int32); s_r2l = raw[4 * nl:4 * (nl + nr)].view(np.int32)
    s_depth = raw[4 * (nl + nr):4 * (2 * nl + nr)].view(np.float32)
    s_p3d = raw[4 * (2 * nl + nr):].view(np.float32).reshape(nl, 3)
    assert (s_l2r == l2r).all() and (s_r2l == r2l).all()
    assert s_depth.tobytes() == depth.tobytes()
    hit = l2r >= 0
    assert s_p3d[hit].tobytes() == p3d[hit].tobytes()
    if mono_l < len(k1) - 100:
        assert hit.sum() > 20 and (l2r[:mono_l] == -1).all() and (l2r[hit] >= mono_r).all()
    else:
        assert hit.sum() <= len(k1) - mono_l


def test_cpp_shim_local_map_two_frames(tmp_path):
    """shim/LocalMap.h (LocalMapMirror::Sync + SearchLocalPoints over cuorb_localmap_*) for two consecutive frames - the
    second with another subset of the map in another order, some points changed in between, another pose - against the
    reference's own Frame::isInFrustum (lifted) followed by the unmodified ORBmatcher::SearchByProjection on the same
    vectors: return value, nToMatch and the final F.mvpMapPoints."""
    from oracle import ref_py as rp
    from tests.matching_cases import extracted_frame, localmap_case
    if not rp.available():
        pytest.skip("oracle/_ref/liborb_ref.so did not travel to this box")
    build.build_libcuorb()
    exe = str(tmp_path / "localmap_check")
    pkg = os.path.join(ROOT, "orbslam3_cpp_fork_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-DCUORB_SHIM_USE_OPENCV",
                           "-I" + os.path.join(ROOT, "oracle", "ref_compat"),
                           os.path.join(ROOT, "tests", "native", "localmap_check.cpp"), "-o", exe,
                           "-L" + pkg, "-l:libcuorb.so", "-Wl,-rpath," + pkg])
    e, k, d = extracted_frame(seed=33)
    sf = np.asarray(e.tables()[0], np.float32)
    w, h, M = 752, 480, 12000
    c1 = localmap_case(k, d, sf, M, 21); c2 = localmap_case(k, d, sf, M, 22)
    rng = np.random.default_rng(4)
    order1 = rng.permutation(M)[:M - 800].astype(np.int32)
    skip1 = rng.permutation(len(order1))[:250].astype(np.int32)
    order2 = rng.permutation(M)[:M - 600].astype(np.int32)
    changed = rng.permutation(M)[:1500].astype(np.int32)
    th, thfar, ratio, mb = 1.0, 30.0, 0.8, 0.1
    keys = ("P", "N", "min_d", "max_d", "desc", "bad", "has_obs")
    v2 = {key: c1[key].copy() for key in keys}
    for key in keys:
        v2[key][changed] = c2[key][changed]

    def expected(data, cam, order, skip_pos):
        f = rp.is_in_frustum(cam, data["P"][order], data["N"][order], data["min_d"][order], data["max_d"][order], 0.5)
        skip = np.zeros(len(order), bool); skip[skip_pos] = True
        in_view = (f["in_view"] > 0) & ~skip & (data["bad"][order] == 0)
        flags = (in_view * 1 + data["bad"][order] * 2 + data["has_obs"][order] * 4).astype(np.uint8)
        fv, keep = op.make_frame_view(k, d, c1["u_right"], c1["bound"], w, h, sf, mb=mb, mbf=c1["mbf"])
        n, best, final = rp.search_by_projection(fv, data["desc"][order], f["proj_x"], f["proj_y"], f["proj_xr"],
                                                 np.maximum(f["level"], 0), f["view_cos"], f["depth"], flags, th, False, thfar, ratio)
        final = np.where(final >= 0, order[np.maximum(final, 0)], final).astype(np.int32)
        return n, int(in_view.sum()), final

    e1 = expected(c1, c1["cam"], order1, skip1)
    e2 = expected(v2, c2["cam"], order2, np.zeros(0, np.int32))
    case = str(tmp_path / "case.bin"); out = str(tmp_path / "out.bin")
    with open(case, "wb") as f:
        np.array([len(k), len(sf), M, len(order1), len(skip1), len(order2), len(changed)], np.int32).tofile(f)
        np.array([th, thfar, ratio, mb, c1["mbf"], w, h], np.float32).tofile(f)
        f.write(bytes(c1["cam"])); f.write(bytes(c2["cam"]))
        np.ascontiguousarray(k).tofile(f); np.ascontiguousarray(d).tofile(f)
        c1["u_right"].tofile(f); c1["bound"].tofile(f); sf.tofile(f)
        for data in (c1, v2):
            for key in keys:
                np.ascontiguousarray(data[key]).tofile(f)
        order1.tofile(f); skip1.tofile(f); order2.tofile(f); changed.tofile(f)
    subprocess.check_call([exe, case, out])
    res = np.fromfile(out, np.int32).reshape(2, 3 + len(k))
    for r, ex in zip(res, (e1, e2)):
        assert int(r[0]) == ex[0] and int(r[1]) == ex[1]
        assert (r[3:] == ex[2]).all()
    assert e1[0] > 300 and e2[0] > 100      # frame 2 looks at mostly frame-1 points from another pose
    in1 = np.zeros(M, bool); in1[order1] = True
    ch = np.zeros(M, bool); ch[changed] = True
    assert int(res[0][2]) == len(order1)
    assert int(res[1][2]) == int((~in1[order2]).sum() + (in1[order2] & ch[order2]).sum())   # only new or changed points were sent


@pytest.mark.parametrize("seed,ori", [(1, True), (2, False)])
def test_cpp_shim_rig_bow(tmp_path, seed, ori):
    """shim/ORBmatcher.h SearchByBoW(KeyFrame*, Frame&, ...) with a rig KeyFrame and a rig Frame against the reference's
    own function (src/ORBmatcher.cc:237-420 with F.Nleft != -1)."""
    from oracle import ref_py as rp
    from tests.bow_cases import make_pair
    from tests.ref_search_cases import bow_rig_views, has_mp_codes
    if not rp.available():
        pytest.skip("oracle/_ref/liborb_ref.so did not travel to this box")
    build.build_libcuorb()
    exe = str(tmp_path / "rig_check")
    pkg = os.path.join(ROOT, "orbslam3_cpp_fork_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-DCUORB_SHIM_USE_OPENCV",
                           "-I" + os.path.join(ROOT, "oracle", "ref_compat"),
                           os.path.join(ROOT, "tests", "native", "rig_check.cpp"), "-o", exe,
                           "-L" + pkg, "-l:libcuorb.so", "-Wl,-rpath," + pkg])
    c = make_pair(seed, tie_heavy=(seed == 2))
    nleft1, nleft2, ratio = 650, 560, 0.75
    h1 = has_mp_codes(c["valid1"], seed)
    (kl, kr, fl, fr), keep = bow_rig_views(c, nleft1, nleft2)
    rn, rml, rmr = rp.search_by_bow_rig(kl, kr, h1, c["fv1"], fl, fr, c["fv2"], ratio, ori)
    n1, n2 = len(c["desc1"]), len(c["desc2"])
    a1, b1, c1_, d1_ = rp.flatten_feature_vector(c["fv1"]); a2, b2, c2_, d2_ = rp.flatten_feature_vector(c["fv2"])
    case = str(tmp_path / "case.bin"); out = str(tmp_path / "out.bin")
    with open(case, "wb") as f:
        np.array([nleft1, n1 - nleft1, nleft2, n2 - nleft2, a1, len(d1_), a2, len(d2_), int(ori)], np.int32).tofile(f)
        np.array([ratio], np.float32).tofile(f)
        for k in keep[0:2]:
            np.ascontiguousarray(k[0]).tofile(f)               # KeyFrame keys: left, right
        np.ascontiguousarray(c["desc1"]).tofile(f); h1.astype(np.uint8).tofile(f)
        for k in keep[2:4]:
            np.ascontiguousarray(k[0]).tofile(f)               # Frame keys: left, right
        np.ascontiguousarray(c["desc2"]).tofile(f)
        b1.tofile(f); c1_.tofile(f); d1_.tofile(f); b2.tofile(f); c2_.tofile(f); d2_.tofile(f)
    subprocess.check_call([exe, "bow", case, out])
    res = np.fromfile(out, np.int32)
    assert int(res[0]) == rn and rn > 80
    assert (res[1:1 + n1] == rml).all() and (res[1 + n1:] == rmr).all()


def test_cpp_shim_local_map_resident_frame(tmp_path):
    """One frame end to end in C++: cuorb_extract leaves the features on the device, then Tracking::SearchLocalPoints
    through shim/LocalMap.h twice - Frame flattened on the host, and SearchLocalPointsResident (keypoints and
    descriptors read in place, Frame::UndistortKeyPoints on the device).  Both against the reference's own
    Frame::isInFrustum (lifted) + unmodified ORBmatcher::SearchByProjection on the host-side undistorted keypoints."""
    from oracle import ref_py as rp
    from orbslam3_cpp_fork_b200 import UndistortKeyPoints, synth
    from tests.matching_cases import extracted_frame, localmap_case
    if not rp.available():
        pytest.skip("oracle/_ref/liborb_ref.so did not travel to this box")
    build.build_libcuorb()
    exe = str(tmp_path / "localmap_resident_check")
    pkg = os.path.join(ROOT, "orbslam3_cpp_fork_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-DCUORB_SHIM_USE_OPENCV",
                           "-I" + os.path.join(ROOT, "oracle", "ref_compat"),
                           os.path.join(ROOT, "tests", "native", "localmap_resident_check.cpp"), "-o", exe,
                           "-L" + pkg, "-l:libcuorb.so", "-Wl,-rpath," + pkg])
    w, h, M, nf = 752, 480, 15000, 1000
    e, k, d = extracted_frame(w, h, nf, seed=35)             # the oracle's extraction = what the device will hold
    sf = np.asarray(e.tables()[0], np.float32)
    K4 = np.array([458.654, 457.296, 367.215, 248.375], np.float32)
    D5 = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0], np.float32)
    kun = UndistortKeyPoints(k, K4, D5)
    bounds = (-25.5, -14.0, w + 30.25, h + 18.0)
    c = localmap_case(kun, d, sf, M, 23)
    rng = np.random.default_rng(6)
    order = rng.permutation(M)[:M - 700].astype(np.int32)
    skip = rng.permutation(len(order))[:200].astype(np.int32)
    th, thfar, ratio, mb = 1.0, 30.0, 0.8, 0.1
    f = rp.is_in_frustum(c["cam"], c["P"][order], c["N"][order], c["min_d"][order], c["max_d"][order], 0.5)
    sk = np.zeros(len(order), bool); sk[skip] = True
    in_view = (f["in_view"] > 0) & ~sk & (c["bad"][order] == 0)
    flags = (in_view * 1 + c["bad"][order] * 2 + c["has_obs"][order] * 4).astype(np.uint8)
    fv, keep = op.make_frame_view(kun, d, np.full(len(k), -1, np.float32), c["bound"], w, h, sf, mb=mb, mbf=c["mbf"],
                                  min_x=bounds[0], min_y=bounds[1], max_x=bounds[2], max_y=bounds[3])
    n_ref, best, final = rp.search_by_projection(fv, c["desc"][order], f["proj_x"], f["proj_y"], f["proj_xr"],
                                                 np.maximum(f["level"], 0), f["view_cos"], f["depth"], flags, th, False, thfar, ratio)
    final = np.where(final >= 0, order[np.maximum(final, 0)], final).astype(np.int32)
    img = str(tmp_path / "img.raw"); case = str(tmp_path / "case.bin"); out = str(tmp_path / "out.bin")
    synth.synth_image(w, h, 35).tofile(img)
    with open(case, "wb") as fh:
        np.array([len(k), len(sf), M, len(order), len(skip), nf], np.int32).tofile(fh)
        np.array([th, thfar, ratio, mb, c["mbf"], *bounds, *K4, *D5], np.float32).tofile(fh)
        fh.write(bytes(c["cam"]))
        c["bound"].tofile(fh); sf.tofile(fh)
        for key in ("P", "N", "min_d", "max_d", "desc", "bad", "has_obs"):
            np.ascontiguousarray(c[key]).tofile(fh)
        order.tofile(fh); skip.tofile(fh)
    subprocess.check_call([exe, img, str(w), str(h), case, out])
    raw = np.fromfile(out, np.uint8)
    N = len(k)
    head = raw[:4 * (1 + 2 * (2 + N))].view(np.int32)
    assert head[0] == N
    assert raw[4 * (1 + 2 * (2 + N)):].tobytes() == kun.tobytes()
    res = head[1:].reshape(2, 2 + N)
    assert n_ref > 200
    for r in res:                                            # host frame, then resident frame
        assert int(r[0]) == n_ref and int(r[1]) == int(in_view.sum())
        assert (r[2:] == final).all()
