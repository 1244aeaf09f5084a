"""GPU (-m gpu): libcuorb through the C ABI against the reference's OWN code (oracle/_ref/liborb_ref.so:
unmodified src/ORBextractor.cc + src/ORBmatcher.cc + Frame::ComputeStereoMatches lifted verbatim,
built in the container by `make -C oracle ref`, shipped prebuilt to the GPU box).  Same bar as
test_gpu_parity.py: bit-exact everywhere, angles included."""
import numpy as np
import pytest

from oracle import oracle_py as op
from oracle import ref_py as rp
from orbslam3_cpp_fork_b200 import synth
from tests.matching_cases import extracted_frame, lastframe_case, mappoint_case
from tests.test_oracle_vs_ref import CASES, replay_mappoints

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not rp.available(), reason="oracle/_ref/liborb_ref.so did not travel to this box")]


@pytest.fixture(scope="module")
def cu():
    import orbslam3_cpp_fork_b200 as c
    if c.load().cuorb_device_count() < 1:
        pytest.fail("no CUDA device: libcuorb has no CPU fallback")
    return c


@pytest.mark.parametrize("w,h,nf,seed,lap,kind", CASES)
def test_extractor_vs_reference(cu, w, h, nf, seed, lap, kind):
    img = synth.degenerate_image(w, h, kind, seed) if kind else synth.synth_image(w, h, seed)
    e = cu.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    r = rp.RefExtractor(nf, 1.2, 8, 20, 7)
    mono, k, d = e(img, None, lap)
    mr, kr, dr = r.extract(img, lap)
    assert (e.GetScaleFactors() == r.tables()[0]).all() and (e.mnFeaturesPerLevel == r.tables()[4]).all()
    for lv in range(8):
        assert (e.download_plane(lv, 0) == r.plane(lv, 0)).all(), f"mvImagePyramid[{lv}] with its border"
    assert mono == mr and len(k) == len(kr)
    assert np.ascontiguousarray(k).tobytes() == kr.tobytes()     # all seven cv::KeyPoint fields, bit for bit
    assert (d == dr).all()


@pytest.mark.parametrize("w,h,nf,seed,fx,mbf", [(752, 480, 1000, 11, 458.654, 47.906), (1241, 376, 2000, 12, 718.856, 386.1448)])
def test_stereo_vs_reference(cu, w, h, nf, seed, fx, mbf):
    L, R = synth.synth_stereo_pair(w, h, seed)
    eL = cu.ORBextractor(nf, max_width=w, max_height=h); eR = cu.ORBextractor(nf, max_width=w, max_height=h)
    eL.extract_batch_async(L[None]); eR.extract_batch_async(R[None])
    cl, _, kl, dl = eL.fetch(); cr, _, kr, dr = eR.fetch()
    mb = mbf / fx
    u, dep = cu.ORBmatcher().SearchStereo(eL, eR, mb, mbf)
    rL, rR = rp.RefExtractor(nf), rp.RefExtractor(nf)
    _, k0, d0 = rL.extract(L); _, k1, d1 = rR.extract(R)
    ur, dr_ = rp.stereo(rL, rR, k0, d0, k1, d1, mb, mbf)
    n = cl[0]
    assert n == len(k0) and (ur >= 0).sum() > 100
    assert u[0, :n].tobytes() == ur.tobytes() and dep[0, :n].tobytes() == dr_.tobytes()


@pytest.mark.parametrize("M,th,with_uright,far", [(50000, 3.0, False, False), (3000, 3.0, True, False)])
def test_search_by_projection_mappoints_vs_reference(cu, M, th, with_uright, far):
    e, k, d = extracted_frame()
    sf = e.tables()[0]
    c = mappoint_case(k, d, sf, M, 77 + M, with_uright=with_uright)
    fv, keep = op.make_frame_view(k, d, c["u_right"], c["bound"], 752, 480, sf, mb=0.1, mbf=40.0)
    n_r, best_r, final_r = rp.search_by_projection(fv, c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"],
                                                   c["depth"], c["flags"], th, far, 60.0, 0.8)
    F = cu.Frame(k, d, sf, 752, 480, mvuRight=c["u_right"], mb=0.1, mbf=40.0)
    F.mvpMapPoints[c["bound"] > 0] = 10 ** 9; F.mvpHasObs[c["bound"] > 0] = 1
    mps = cu.MapPointSoA(c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"], c["depth"],
                         (c["flags"] & 1) > 0, (c["flags"] & 2) > 0, (c["flags"] & 4) > 0)
    n, best = cu.ORBmatcher(0.8).SearchByProjection(F, mps, th, far, 60.0)
    assert n == n_r and (best == best_r).all()
    assert (replay_mappoints(best, c["bound"], len(k)) == final_r).all()


@pytest.mark.parametrize("fwd,bwd,th", [(False, False, 7.0), (True, False, 15.0), (False, True, 7.0)])
def test_search_by_projection_lastframe_vs_reference(cu, fwd, bwd, th):
    e, k, d = extracted_frame(seed=33)
    sf = e.tables()[0]
    c = lastframe_case(k, d, 5)
    z = (np.float32(1.0) / c["inv_z"]).astype(np.float32)
    inv_z = (1.0 / z.astype(np.float64)).astype(np.float32)
    fv, keep = op.make_frame_view(k, d, c["u_right"], None, 752, 480, sf, mb=0.1, mbf=40.0)
    n_r, final_r = rp.search_by_projection_frame(fv, c["desc"], c["u"], c["v"], z, c["octave"], c["angle"], c["flags"],
                                                 th, fwd, bwd, True)
    F = cu.Frame(k, d, sf, 752, 480, mvuRight=c["u_right"], mb=0.1, mbf=40.0)
    n, best, rot, keepb = cu.ORBmatcher(0.9, True).SearchByProjectionLastFrame(
        F, c["desc"], c["u"], c["v"], inv_z, c["octave"], c["angle"], c["flags"], th, fwd, bwd)
    final = np.full(len(k), -1, np.int32)
    for i, b in enumerate(best):
        if b >= 0:
            final[b] = i
    for i, b in enumerate(best):
        if b >= 0 and rot[i] not in keepb:
            final[b] = -1
    assert n == n_r and (final == final_r).all()
