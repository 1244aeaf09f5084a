"""GPU (-m gpu): cuorb_localmap_search_resident - the local-map search on a Frame whose keypoints, descriptors and
uRight never left the device (Frame::UndistortKeyPoints + AssignFeaturesToGrid + the flattening
for the search done where the extractor wrote its output).  The bar is identity with cuorb_localmap_search on the same
Frame built on the host (which tests/test_gpu_localmap.py pins to the oracle and to the reference's own functions), and
with the oracle directly for one case."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle_py as op
from tests.matching_cases import localmap_case, localmap_expected

pytestmark = pytest.mark.gpu
W, H = 752, 480
K4 = np.array([458.654, 457.296, 367.215, 248.375], np.float32)          # EuRoC cam0
D5 = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0], np.float32)


@pytest.fixture(scope="module")
def batch():
    """Two images through the GPU extractor; the handle keeps their features on the device."""
    from orbslam3_cpp_fork_b200 import ORBextractor, synth
    e = ORBextractor(1000, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=2)
    e.extract_batch_async(np.stack([synth.synth_image(W, H, 33), synth.synth_image(W, H, 34)]))
    cnt, mono, k, d = e.fetch()
    return e, cnt, k, d, np.asarray(e.GetScaleFactors(), np.float32)


def _cam(c):
    from orbslam3_cpp_fork_b200 import _lib
    return _lib.FrustumCamera.from_buffer_copy(bytes(c["cam"]))


def _frame(k, d, sf, c, u_right=None, **kw):
    from orbslam3_cpp_fork_b200.matcher import Frame
    F = Frame(k, d, sf, W, H, mvuRight=u_right, mb=0.1, mbf=c["mbf"], **kw)
    F.mvpMapPoints[c["bound"] > 0] = 10 ** 6; F.mvpHasObs[c["bound"] > 0] = 1
    return F


def _map(c, extra=0):
    from orbslam3_cpp_fork_b200 import LocalMap
    m = LocalMap(len(c["P"]) + extra)
    m.update(0, c["P"], c["N"], c["min_d"], c["max_d"], c["desc"], c["bad"], c["has_obs"])
    return m


def _same(a, b):
    assert a[0] == b[0] and a[1] == b[1], (a[:2], b[:2])
    for x, y in zip(a[2:5], b[2:5]):
        assert (x == y).all()


@pytest.mark.parametrize("image_index,M,seed,th,far", [(0, 30000, 11, 1.0, False), (1, 12000, 12, 3.0, True)])
def test_resident_equals_host_frame(batch, image_index, M, seed, th, far):
    e, cnt, ks, ds, sf = batch
    n = int(cnt[image_index]); k, d = ks[image_index, :n], ds[image_index, :n]
    c = localmap_case(k, d, sf, M, seed)
    m = _map(c)
    host = m.search(_cam(c), _frame(k, d, sf, c), c["skip_idx"], 0.5, th, far, 30.0, 0.8)
    Fr = _frame(k, d, sf, c)
    Fr.mvKeysUn = np.zeros_like(Fr.mvKeysUn); Fr.mDescriptors = np.zeros_like(Fr.mDescriptors)   # must not be read
    res = m.search_extracted(_cam(c), Fr, e, image_index, skip_idx=c["skip_idx"], th=th, bFarPoints=far, thFarPoints=30.0)
    _same(host, res)
    assert res[0] > 300
    if image_index == 0:     # and the oracle, made of its two parts, with a monocular frame (uRight = -1 everywhere)
        cm = dict(c); cm["u_right"] = np.full(n, -1, np.float32)
        o_in, o_n, o_best = localmap_expected(cm, k, d, sf, op.is_in_frustum, op.search_by_projection, th, far, 30.0, 0.8)
        assert res[1] == o_in and res[0] == o_n and (res[4] == o_best).all()


def test_resident_undistorts_on_device(batch):
    """mvKeys stay on the device; UndistortKeyPoints runs there and mvKeysUn comes back with the matches."""
    from orbslam3_cpp_fork_b200 import UndistortKeyPoints
    e, cnt, ks, ds, sf = batch
    n = int(cnt[0]); k, d = ks[0, :n], ds[0, :n]
    kun = UndistortKeyPoints(k, K4, D5)
    assert (kun["x"] != k["x"]).any()
    bounds = dict(mnMinX=-22.5, mnMinY=-13.25, mnMaxX=W + 31.0, mnMaxY=H + 17.5)   # ComputeImageBounds of a distorted image
    c = localmap_case(kun, d, sf, 20000, 13)
    m = _map(c)
    host = m.search(_cam(c), _frame(kun, d, sf, c, **bounds), c["skip_idx"], 0.5, 1.0, False, 30.0, 0.8)
    res = m.search_extracted(_cam(c), _frame(k, d, sf, c, **bounds), e, 0, K4=K4, dist5=D5, want_keys_un=True,
                             skip_idx=c["skip_idx"])
    _same(host, res)
    assert res[-1].tobytes() == kun.tobytes()
    assert res[0] > 300
    # mDistCoef[0] == 0: the reference copies mvKeys (src/Frame.cc:888-892), the remaining coefficients are not applied
    d0 = D5.copy(); d0[0] = 0
    res0 = m.search_extracted(_cam(c), _frame(k, d, sf, c, **bounds), e, 0, K4=K4, dist5=d0, want_keys_un=True,
                              skip_idx=c["skip_idx"])
    host0 = m.search(_cam(c), _frame(k, d, sf, c, **bounds), c["skip_idx"], 0.5, 1.0, False, 30.0, 0.8)
    _same(host0, res0)
    assert res0[-1].tobytes() == k.tobytes()


def test_resident_stereo_uright_read_in_place():
    """Stereo: uRight is read from the left handle where cuorb_stereo_match left it."""
    from orbslam3_cpp_fork_b200 import ORBextractor, ORBmatcher, synth
    left, right = synth.synth_stereo_pair(W, H, seed=9)
    eL = ORBextractor(1200, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1)
    eR = ORBextractor(1200, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=1)
    eL.extract_batch_async(left[None]); eR.extract_batch_async(right[None])
    cnt, _, ks, ds = eL.fetch(); eR.fetch()
    mbf = 47.906
    u, _ = ORBmatcher().SearchStereo(eL, eR, 0.11, mbf)
    n = int(cnt[0]); k, d, ur = ks[0, :n], ds[0, :n], u[0, :n].copy()
    assert (ur > 0).sum() > 100
    sf = np.asarray(eL.GetScaleFactors(), np.float32)
    c = localmap_case(k, d, sf, 25000, 14)
    c["mbf"] = mbf
    m = _map(c)
    host = m.search(_cam(c), _frame(k, d, sf, c, u_right=ur), c["skip_idx"], 0.5, 1.0, True, 30.0, 0.8)
    res = m.search_extracted(_cam(c), _frame(k, d, sf, c, u_right=np.zeros(n, np.float32)), eL, 0, skip_idx=c["skip_idx"],
                             bFarPoints=True, thFarPoints=30.0)
    _same(host, res)
    mono = m.search(_cam(c), _frame(k, d, sf, c), c["skip_idx"], 0.5, 1.0, True, 30.0, 0.8)
    assert (mono[4] != host[4]).any()            # the right-coordinate test did reject something


def test_resident_cell_order_decides_ties():
    """Many features with the same descriptor a pixel apart: equally distant candidates everywhere, so the outcome
    hangs on the order of the features inside a grid cell (ascending index, as mGrid's vectors hold them)."""
    import torch
    from orbslam3_cpp_fork_b200 import _lib
    rng = np.random.default_rng(5)
    n = 6000
    k = np.zeros(n, _lib.KP_DTYPE)
    k["x"] = rng.integers(40, 200, n) + rng.integers(0, 4, n) * 0.25        # ~160 x 120 px: tens of features per cell
    k["y"] = rng.integers(40, 160, n) + rng.integers(0, 4, n) * 0.25
    k["octave"] = rng.integers(0, 3, n); k["size"] = 31; k["angle"] = 10; k["response"] = 50; k["class_id"] = -1
    proto = rng.integers(0, 256, (8, 32), dtype=np.uint8)
    d = proto[rng.integers(0, 8, n)]                                        # 8 distinct descriptors in the whole frame
    sf = (1.2 ** np.arange(8)).astype(np.float32)
    c = localmap_case(k, d, sf, 10000, 15)
    m = _map(c)
    host = m.search(_cam(c), _frame(k, d, sf, c), [], 0.5, 3.0, False, 30.0, 1.01)
    dk = torch.from_numpy(k.view(np.uint8).copy()).cuda(); dd = torch.from_numpy(d.copy()).cuda()
    torch.cuda.synchronize()
    for _ in range(3):
        res = m.search_resident(_cam(c), _frame(k, d, sf, c), dk.data_ptr(), dd.data_ptr(), th=3.0, thFarPoints=30.0, nnratio=1.01)
        _same(host, res)
    assert host[0] > 50


def test_resident_rejects_host_pointers_and_empty_frame(batch):
    from orbslam3_cpp_fork_b200 import _lib
    e, cnt, ks, ds, sf = batch
    n = int(cnt[0]); k, d = ks[0, :n], ds[0, :n]
    c = localmap_case(k, d, sf, 3000, 16)
    m = _map(c)
    F = _frame(k, d, sf, c)
    with pytest.raises(_lib.CuorbError):
        m.search_resident(_cam(c), F, F.mvKeysUn.ctypes.data, F.mDescriptors.ctypes.data)
    assert "device pointers" in _lib.last_error()
    with pytest.raises(_lib.CuorbError):                       # K4 without dist5
        m.search_extracted(_cam(c), F, e, 0, K4=K4)
    F0 = _frame(k[:0], d[:0], sf, dict(c, bound=c["bound"][:0]))
    r = m.search_extracted(_cam(c), F0, e, 0)
    assert r[0] == 0 and r[1] > 0 and (r[4] == -1).all()
    with pytest.raises(_lib.CuorbError):
        m.search_extracted(_cam(c), F, e, 5)                   # no such image in the last batch
    dk, dd = C.c_void_p(), C.c_void_p()
    cap = _lib.load().cuorb_extractor_device_features(e.handle, 1, C.byref(dk), C.byref(dd), None, None)
    assert cap == e.capacity and dk.value and dd.value
