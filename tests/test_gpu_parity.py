"""GPU (-m gpu): libcuorb through the C ABI vs the CPU oracle and the committed golden vectors.
Bar: bit-exact pyramid, blur, FAST candidates, keypoints (order, octave, response, size),
descriptors, Hamming distances, match indices, SAD, mvuRight/mvDepth; angles within 1e-4 rad
(expected and asserted bit-exact on these inputs)."""
import glob
import os

import numpy as np
import pytest

from oracle import oracle_py as op
from orbslam3_cpp_fork_b200 import synth
from tests.matching_cases import extracted_frame, lastframe_case, mappoint_case
from tests.test_oracle_golden import GOLDEN, golden_image

pytestmark = pytest.mark.gpu

ANGLE_TOL_DEG = np.degrees(1e-4)


@pytest.fixture(scope="module")
def cu():
    import orbslam3_cpp_fork_b200 as c
    if c.load().cuorb_device_count() < 1:
        pytest.fail("no CUDA device: libcuorb has no CPU fallback")
    return c


def assert_extract_equal(k, d, ko, do):
    assert len(k) == len(ko)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert (k[f] == ko[f]).all(), f
    assert np.abs(k["angle"] - ko["angle"]).max(initial=0) <= ANGLE_TOL_DEG
    assert (k["angle"] == ko["angle"]).all()
    assert (d == do).all()


CASES = [(752, 480, 1000, 1, (0, 0), None), (752, 480, 1000, 2, (0, 1000), None), (640, 480, 1000, 3, (0, 0), None),
         (1241, 376, 2000, 4, (0, 0), None), (320, 240, 500, 5, (100, 200), None), (752, 480, 5000, 8, (0, 0), None),
         (752, 480, 1000, 6, (0, 0), "black"), (752, 480, 1000, 6, (0, 0), "flat"), (752, 480, 1000, 6, (0, 0), "noise")]


@pytest.mark.parametrize("w,h,nf,seed,lap,kind", CASES)
def test_extractor_stage_parity(cu, w, h, nf, seed, lap, kind):
    img = synth.degenerate_image(w, h, kind, seed) if kind else synth.synth_image(w, h, seed)
    e = cu.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    o = op.OracleExtractor(nf, 1.2, 8, 20, 7)
    mono, k, d = e(img, None, lap)
    mo, ko, do = o.extract(img, lap)
    assert (e.GetScaleFactors() == o.tables()[0]).all() and (e.mnFeaturesPerLevel == o.tables()[4]).all()
    for lv in range(8):
        assert (e.download_plane(lv, 0) == o.plane(lv, 0)).all(), f"pyramid level {lv}"
        ob = o.plane(lv, 2)
        if ob is not None:
            assert (e.download_plane(lv, 2) == ob).all(), f"blur level {lv}"
        xs, ys, rs = e.download_candidates(lv)
        xo, yo, ro = o.candidates(lv)
        assert len(xs) == len(xo) and (xs == xo).all() and (ys == yo).all() and (rs == ro).all(), f"FAST level {lv}"
    assert mono == mo
    assert_extract_equal(k, d, ko, do)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_golden_vectors(cu, path):
    g = np.load(path)
    img = golden_image(g)
    h, w = img.shape
    e = cu.ORBextractor(int(g["nfeatures"]), 1.2, 8, 20, 7, max_width=w, max_height=h)
    mono, k, d = e(img, None, tuple(int(v) for v in g["lap"]))
    assert mono == int(g["mono"])
    assert_extract_equal(k, d, g["kps"], g["desc"])


def test_empty_and_oversize(cu):
    e = cu.ORBextractor(500, max_width=320, max_height=240)
    mono, k, d = e(np.zeros((0, 0), np.uint8))
    assert mono == -1 and len(k) == 0 and d.shape == (0, 32)            # src/ORBextractor.cc:1060
    L = cu.load()
    buf = np.zeros(4, np.uint8)
    assert L.cuorb_extract(e.handle, None, 0, 0, 0, 0, 0, None, None, 0, None) == -1
    with pytest.raises(cu.CuorbError):
        e(np.zeros((480, 752), np.uint8))                                # larger than the arena
    with pytest.raises(cu.CuorbError):
        cu.ORBextractor(500, max_width=60, max_height=60)                # levels smaller than a FAST cell


def test_ragged_batch_and_size_change(cu):
    """A batch mixing textured, black, flat and noise frames == per-image results; a handle reused
    for a smaller image size rebuilds its geometry."""
    w, h = 640, 480
    imgs = np.stack([synth.synth_image(w, h, 40), synth.degenerate_image(w, h, "black"), synth.synth_image(w, h, 41),
                     synth.degenerate_image(w, h, "flat", 2), synth.degenerate_image(w, h, "noise", 3)])
    e = cu.ORBextractor(1000, max_width=w, max_height=h, max_batch=8)
    e.extract_batch_async(imgs)
    counts, mono, kps, desc = e.fetch()
    o = op.OracleExtractor(1000)
    for i in range(len(imgs)):
        mo, ko, do = o.extract(imgs[i])
        assert counts[i] == len(ko) and mono[i] == mo
        assert_extract_equal(kps[i, :counts[i]], desc[i, :counts[i]], ko, do)
    small = synth.synth_image(400, 300, 42)
    mo, ko, do = o.extract(small)
    m2, k2, d2 = e(small)
    assert m2 == mo
    assert_extract_equal(k2, d2, ko, do)


def test_large_batch_idempotent(cu):
    """Full-size property (BASELINE cfg2 shape): 64 frames, every copy of a frame gives the same
    result as its first occurrence, and running the batch twice is idempotent."""
    w, h = 640, 480
    base = [synth.synth_image(w, h, 60 + i) for i in range(4)]
    imgs = np.stack([base[i % 4] for i in range(64)])
    e = cu.ORBextractor(1000, max_width=w, max_height=h, max_batch=64)
    e.extract_batch_async(imgs); c1, m1, k1, d1 = e.fetch()
    e.extract_batch_async(imgs); c2, m2, k2, d2 = e.fetch()
    assert (c1 == c2).all() and (k1 == k2).all() and (d1 == d2).all()
    for i in range(4, 64):
        assert c1[i] == c1[i % 4] and (k1[i] == k1[i % 4]).all() and (d1[i] == d1[i % 4]).all()
    o = op.OracleExtractor(1000)
    _, ko, do = o.extract(base[3])
    assert_extract_equal(k1[63, :c1[63]], d1[63, :c1[63]], ko, do)


def test_stereo_match_then_next_extraction_is_ordered(cu):
    """The async stereo match reads the RIGHT handle's pyramid / keypoints / descriptors on the LEFT
    handle's stream; an extraction enqueued on the right handle straight afterwards (no host sync)
    must not overwrite them under the match (cuorb.h: cuorb_stereo_match_batch_async)."""
    w, h, nb = 752, 480, 32
    pairs = [synth.synth_stereo_pair(w, h, 300 + i) for i in range(4)]
    Ls = np.stack([pairs[i % 4][0] for i in range(nb)]); Rs = np.stack([pairs[i % 4][1] for i in range(nb)])
    other = np.stack([synth.synth_image(w, h, 900 + (i % 3)) for i in range(nb)])
    eL = cu.ORBextractor(1000, max_width=w, max_height=h, max_batch=nb)
    eR = cu.ORBextractor(1000, max_width=w, max_height=h, max_batch=nb)
    m = cu.ORBmatcher()
    mb, mbf = 47.906 / 458.654, 47.906
    eL.extract_batch_async(Ls); eR.extract_batch_async(Rs)
    eL.sync(); eR.sync()
    u0, d0 = m.SearchStereo(eL, eR, mb, mbf)                       # reference run, fully synchronised
    from orbslam3_cpp_fork_b200 import _lib
    L_ = _lib.load()
    for _ in range(5):
        eL.extract_batch_async(Ls); eR.extract_batch_async(Rs)
        _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, mb, mbf))
        eR.extract_batch_async(other)                              # would race with the match if unordered
        u = np.zeros_like(u0); d = np.zeros_like(d0)
        _lib.check(L_.cuorb_stereo_fetch(eL.handle, _lib.ptr(u), _lib.ptr(d), eL.capacity, None, None, None))
        eR.sync()
        assert (u == u0).all() and (d == d0).all()


@pytest.mark.parametrize("w,h,nf,seed,fx,mbf", [(752, 480, 1000, 11, 458.654, 47.906), (1241, 376, 2000, 12, 718.856, 386.1448)])
def test_stereo_parity(cu, w, h, nf, seed, fx, mbf):
    L, R = synth.synth_stereo_pair(w, h, seed)
    eL = cu.ORBextractor(nf, max_width=w, max_height=h, max_batch=2)
    eR = cu.ORBextractor(nf, max_width=w, max_height=h, max_batch=2)
    black = np.zeros_like(L)
    eL.extract_batch_async(np.stack([L, black])); eR.extract_batch_async(np.stack([R, R]))
    cl, _, kl, dl = eL.fetch(); cr, _, kr, dr = eR.fetch()
    mb = mbf / fx
    m = cu.ORBmatcher()
    u, dep, dbg = m.SearchStereo(eL, eR, mb, mbf, debug=True)
    oL, oR = op.OracleExtractor(nf), op.OracleExtractor(nf)
    _, k0, d0 = oL.extract(L); _, k1, d1 = oR.extract(R)
    assert_extract_equal(kl[0, :cl[0]], dl[0, :cl[0]], k0, d0)
    so = op.stereo(oL, oR, k0, d0, k1, d1, mb, mbf)
    n = cl[0]
    assert (so["u_right"] >= 0).sum() > 100
    assert (dbg["best_idx_r"][0, :n] == so["best_idx_r"]).all() and (dbg["best_dist"][0, :n] == so["best_dist"]).all()
    assert (dbg["sad"][0, :n] == so["sad"]).all()
    assert (u[0, :n] == so["u_right"]).all() and (dep[0, :n] == so["depth"]).all()   # bit-exact floats
    assert cl[1] == 0                                                                # black left image: nothing to match
    # host-pointer form (the member set Frame::ComputeStereoMatches reads)
    u2, d2 = m.ComputeStereoMatches(eL, eR, k0, d0, k1, d1, mb, mbf)
    assert (u2 == so["u_right"]).all() and (d2 == so["depth"]).all()


def test_hamming_batch(cu):
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, (3000, 32), dtype=np.uint8); b = rng.integers(0, 256, (2000, 32), dtype=np.uint8)
    b[:100] = a[:100]
    ia = rng.integers(0, 3000, 50000).astype(np.int32); ib = rng.integers(0, 2000, 50000).astype(np.int32)
    out = cu.ORBmatcher().DescriptorDistanceBatch(a, b, ia, ib)
    ref = np.unpackbits(a[ia] ^ b[ib], axis=1).sum(1)
    assert (out == ref).all()
    assert (cu.ORBmatcher().DescriptorDistanceBatch(a[:100], b[:100]) == 0).all()
    assert cu.ORBmatcher.DescriptorDistance(a[5], b[7]) == op.hamming(a[5], b[7])


@pytest.mark.parametrize("M,th,with_uright,far", [(50000, 3.0, False, False), (50000, 1.0, True, True), (3000, 3.0, True, False)])
def test_search_by_projection_mappoints(cu, M, th, with_uright, far):
    """cfg4: 1000 frame descriptors vs M local-map MapPoints (order-dependent binding included)."""
    e, k, d = extracted_frame()
    sf = e.tables()[0]
    c = mappoint_case(k, d, sf, M, 77 + M, with_uright=with_uright)
    fv, keep = op.make_frame_view(k, d, c["u_right"], c["bound"], 752, 480, sf, mb=0.1, mbf=40.0)
    n_o, best_o = op.search_by_projection(fv, c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"],
                                          c["depth"], c["flags"], th, far, 60.0, 0.8)
    F = cu.Frame(k, d, sf, 752, 480, mvuRight=c["u_right"], mb=0.1, mbf=40.0)
    F.mvpMapPoints[c["bound"] > 0] = 10 ** 9; F.mvpHasObs[c["bound"] > 0] = 1
    mps = cu.MapPointSoA(c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"], c["depth"],
                         (c["flags"] & 1) > 0, (c["flags"] & 2) > 0, (c["flags"] & 4) > 0)
    n, best = cu.ORBmatcher(0.8).SearchByProjection(F, mps, th, far, 60.0)
    assert n_o > 100
    assert n == n_o and (best == best_o).all()


@pytest.mark.parametrize("fwd,bwd,th", [(False, False, 7.0), (True, False, 15.0), (False, True, 7.0)])
def test_search_by_projection_lastframe(cu, fwd, bwd, th):
    e, k, d = extracted_frame(seed=33)
    sf = e.tables()[0]
    c = lastframe_case(k, d, 5)
    fv, keep = op.make_frame_view(k, d, c["u_right"], None, 752, 480, sf, mb=0.1, mbf=40.0)
    n_o, best_o, rot_o, keep_o = op.search_by_projection_frame(fv, c["desc"], c["u"], c["v"], c["inv_z"], c["octave"],
                                                              c["angle"], c["flags"], th, fwd, bwd, True)
    F = cu.Frame(k, d, sf, 752, 480, mvuRight=c["u_right"], mb=0.1, mbf=40.0)
    n, best, rot, keepb = cu.ORBmatcher(0.9, True).SearchByProjectionLastFrame(
        F, c["desc"], c["u"], c["v"], c["inv_z"], c["octave"], c["angle"], c["flags"], th, fwd, bwd)
    assert n_o > 50
    assert n == n_o and (best == best_o).all() and (rot == rot_o).all() and (keepb == keep_o).all()


@pytest.mark.parametrize("w,h,nf,scale,nlevels,ini,mn,seed", [
    (641, 479, 1000, 1.2, 8, 20, 7, 91),       # odd size: unaligned pitches, ragged last cells
    (1280, 720, 1500, 1.2, 8, 20, 7, 92),
    (1920, 1080, 2000, 1.2, 8, 20, 7, 93),     # largest size the shim defaults to
    (752, 480, 1000, 1.1, 12, 20, 7, 94),      # many shallow levels
    (752, 480, 800, 1.5, 5, 20, 7, 95),        # steep pyramid: wide resize windows
    (400, 300, 300, 1.3, 6, 12, 12, 96),       # iniThFAST == minThFAST: no retry pass
])
def test_other_geometries(cu, w, h, nf, scale, nlevels, ini, mn, seed):
    """Sizes, pyramid shapes and thresholds other than the benchmark's: keypoints, descriptors and the
    per-level FAST candidates stay bit-identical to the oracle."""
    img = synth.synth_image(w, h, seed)
    e = cu.ORBextractor(nf, scale, nlevels, ini, mn, max_width=w, max_height=h)
    o = op.OracleExtractor(nf, scale, nlevels, ini, mn)
    mono, k, d = e(img, None, (0, 0))
    mo, ko, do = o.extract(img, (0, 0))
    for lv in range(nlevels):
        assert (e.download_plane(lv, 0) == o.plane(lv, 0)).all(), f"pyramid level {lv}"
        xs, ys, rs = e.download_candidates(lv)
        xo, yo, ro = o.candidates(lv)
        assert len(xs) == len(xo) and (xs == xo).all() and (ys == yo).all() and (rs == ro).all(), f"FAST level {lv}"
    assert mono == mo
    assert_extract_equal(k, d, ko, do)


VARIANTS = [{"CUORB_FAST_LEGACY": "1"}, {"CUORB_SIDE_BLUR": "0"}, {"CUORB_FAST_CTAS": "0"}, {"CUORB_FAST_CTAS": "3"}, {"CUORB_PDL": "0"}]


@pytest.mark.parametrize("env_over", VARIANTS, ids=lambda d: ",".join(f"{k[6:]}={v}" for k, v in d.items()))
def test_kernel_variants_subprocess(cu, env_over):
    """The four test / tuning hooks of DESIGN.md section 6 select launch shapes or the wide-cell kernel, which must
    stay bit-identical to the oracle (one-CTA-per-cell FAST on ordinary geometry, serial blur, other persistent-grid
    sizes, plain stream order instead of dependent launches for this two-image batch).  The hooks are read once per process, hence the subprocess."""
    import subprocess, sys, textwrap
    code = textwrap.dedent("""
        import numpy as np
        import orbslam3_cpp_fork_b200 as c
        from oracle import oracle_py as op
        from orbslam3_cpp_fork_b200 import synth
        for (w, h, nf, seed) in ((752, 480, 1000, 21), (1241, 376, 2000, 22), (333, 251, 300, 23)):
            img = synth.synth_image(w, h, seed)
            e = c.ORBextractor(nf, max_width=w, max_height=h, max_batch=2)
            e.extract_batch_async(np.stack([img, img[::-1].copy()]))
            cnt, mono, k, d = e.fetch()
            o = op.OracleExtractor(nf)
            _, ko, do = o.extract(img)
            assert cnt[0] == len(ko) and (d[0, :cnt[0]] == do).all()
            for f in ("x", "y", "angle", "response", "octave"):
                assert (k[0, :cnt[0]][f] == ko[f]).all(), f
        print("variant ok")
    """)
    env = dict(os.environ, **env_over)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], env=env, cwd=root, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "variant ok" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("w,h,nf,seed,fx,mbf", [(752, 480, 1000, 11, 458.654, 47.906), (1241, 376, 2000, 12, 718.856, 386.1448)])
def test_fused_stereo_frame(cu, w, h, nf, seed, fx, mbf):
    """cuorb_extract_stereo (one call = ExtractORB x 2 + ComputeStereoMatches) against the oracle, twice in a
    row on one handle pair (the packed result buffer is reused), then a black right image (no right keypoints)."""
    L, R = synth.synth_stereo_pair(w, h, seed)
    eL = cu.ORBextractor(nf, max_width=w, max_height=h, max_batch=1)
    eR = cu.ORBextractor(nf, max_width=w, max_height=h, max_batch=1)
    mb = mbf / fx
    oL, oR = op.OracleExtractor(nf), op.OracleExtractor(nf)
    m = cu.ORBmatcher()
    for rep, (a, b) in enumerate(((L, R), (R, L), (L, R))):
        mono, kl, dl, kr, dr, u, d = m.ExtractStereo(eL, eR, a, b, mb, mbf)
        m0, k0, d0 = oL.extract(a); m1, k1, d1 = oR.extract(b)
        assert mono == m0
        assert_extract_equal(kl, dl, k0, d0)
        assert_extract_equal(kr, dr, k1, d1)
        so = op.stereo(oL, oR, k0, d0, k1, d1, mb, mbf)
        assert u.tobytes() == so["u_right"].tobytes() and d.tobytes() == so["depth"].tobytes(), rep
    assert (u >= 0).sum() > 100
    mono, kl, dl, kr, dr, u, d = m.ExtractStereo(eL, eR, L, np.zeros_like(L), mb, mbf)
    assert len(kr) == 0 and len(kl) == len(k0) and (u == -1).all() and (d == -1).all()
    # handles made for batches cannot use the fused call
    eB = cu.ORBextractor(nf, max_width=w, max_height=h, max_batch=2)
    with pytest.raises(cu.CuorbError):
        m.ExtractStereo(eB, eR, L, R, mb, mbf)


def test_single_image_handle_packed_fetch(cu):
    """cuorb_extract on a max_batch == 1 handle fetches through one packed copy: same results as a batch handle,
    counts still written and CUORB_E_CAPACITY returned when the caller's arrays are too small, empty frame -> n = 0."""
    import ctypes as C
    from orbslam3_cpp_fork_b200 import _lib
    w, h = 640, 480
    img = synth.synth_image(w, h, 51)
    e1 = cu.ORBextractor(1000, max_width=w, max_height=h, max_batch=1)
    e4 = cu.ORBextractor(1000, max_width=w, max_height=h, max_batch=4)
    a = e1(img, None, (100, 300)); b = e4(img, None, (100, 300))
    assert a[0] == b[0] and a[1].tobytes() == b[1].tobytes() and (a[2] == b[2]).all()
    mo, ko, do = op.OracleExtractor(1000).extract(img, (100, 300))
    assert a[0] == mo
    assert_extract_equal(a[1], a[2], ko, do)
    L_ = cu.load()
    small = 100
    kps = np.zeros(small, _lib.KP_DTYPE); desc = np.zeros((small, 32), np.uint8); n = C.c_int(0)
    rc = L_.cuorb_extract(e1.handle, _lib.ptr(img), w, h, w, 0, 0, _lib.ptr(kps), _lib.ptr(desc), small, C.byref(n))
    assert rc == _lib.E_CAPACITY and n.value == len(ko)
    mono, k, d = e1(np.zeros((h, w), np.uint8))
    assert mono == 0 and len(k) == 0 and d.shape == (0, 32)
    a2 = e1(img, None, (100, 300))
    assert a2[1].tobytes() == a[1].tobytes() and (a2[2] == a[2]).all()
