"""Host-side mirror of the per-frame glue the reference runs right after ORBextractor::operator()
(SURVEY.md §8f row N2): Frame::UndistortKeyPoints (src/Frame.cc:886-913) and
Frame::ComputeStereoFromRGBD (src/Frame.cc:1115-1133).  Compute is libcuorb's (no CPU fallback).
"""
from __future__ import annotations

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, check, ptr
from .extractor import _as_device_ptr


def UndistortKeyPoints(mvKeys, K, mDistCoef, device=0):
    """mvKeys: KP_DTYPE array; K: 3x3 camera matrix or (fx, fy, cx, cy); mDistCoef: k1, k2, p1, p2[, k3].
    Returns mvKeysUn (a copy of mvKeys when mDistCoef[0] == 0, as the reference does)."""
    L = _lib.load()
    kps = np.ascontiguousarray(mvKeys, dtype=KP_DTYPE)
    K = np.asarray(K, dtype=np.float32)
    K4 = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], np.float32) if K.ndim == 2 else np.ascontiguousarray(K, np.float32)
    d5 = np.zeros(5, np.float32)
    d = np.asarray(mDistCoef, np.float32).ravel()
    d5[:min(5, len(d))] = d[:5]
    out = np.empty_like(kps)
    check(L.cuorb_undistort_keypoints(int(device), len(kps), ptr(kps), ptr(K4), ptr(d5), ptr(out)))
    return out


def ComputeStereoFromRGBD(mvKeys, mvKeysUn, imDepth, mbf, device=0):
    """imDepth: float32 [h, w] numpy (host) or torch CUDA tensor.  Returns (mvuRight, mvDepth)."""
    L = _lib.load()
    kps = np.ascontiguousarray(mvKeys, dtype=KP_DTYPE)
    kun = np.ascontiguousarray(mvKeysUn, dtype=KP_DTYPE)
    dev = _as_device_ptr(imDepth)
    if dev is not None:
        p, shape, strides = dev
        where = _lib.MEM_DEVICE
    else:
        imDepth = np.asarray(imDepth)
        if imDepth.dtype != np.float32:
            raise ValueError("imDepth must be float32 (CV_32F)")
        p, shape, strides, where = imDepth.ctypes.data, imDepth.shape, imDepth.strides, _lib.MEM_HOST
    if len(shape) != 2 or strides[1] != 4:
        raise ValueError("imDepth must be [h, w] float32 with unit pixel stride")
    u = np.empty(len(kps), np.float32)
    z = np.empty(len(kps), np.float32)
    check(L.cuorb_stereo_from_rgbd(int(device), len(kps), ptr(kps), ptr(kun), ptr(p), shape[1], shape[0], strides[0],
                                   where, float(mbf), ptr(u), ptr(z)))
    return u, z


def isInFrustum(camera, world_pos, normal, mfMinDistance, mfMaxDistance, viewingCosLimit, device=0):
    """Frame::isInFrustum (src/Frame.cc:644-707) for n MapPoints at once.  camera: _lib.FrustumCamera (mRcw, mtcw,
    mOw, fx, fy, cx, cy, mbf, image bounds, mfLogScaleFactor, mnScaleLevels).  Returns a dict with mbTrackInView,
    mTrackProjX, mTrackProjY, mTrackProjXR, mnTrackScaleLevel, mTrackViewCos, mTrackDepth (numpy arrays) and
    the number of points in view — exactly the fields ORBmatcher::SearchByProjection(F, vpMapPoints) reads."""
    import ctypes as C
    L = _lib.load()
    P = np.ascontiguousarray(world_pos, np.float32).reshape(-1, 3)
    N = np.ascontiguousarray(normal, np.float32).reshape(-1, 3)
    mn = np.ascontiguousarray(mfMinDistance, np.float32); mx = np.ascontiguousarray(mfMaxDistance, np.float32)
    n = len(P)
    out = dict(mbTrackInView=np.zeros(n, np.uint8), mTrackProjX=np.zeros(n, np.float32), mTrackProjY=np.zeros(n, np.float32),
               mTrackProjXR=np.zeros(n, np.float32), mnTrackScaleLevel=np.zeros(n, np.int32),
               mTrackViewCos=np.zeros(n, np.float32), mTrackDepth=np.zeros(n, np.float32))
    out["count"] = check(L.cuorb_is_in_frustum(int(device), C.byref(camera), n, ptr(P), ptr(N), ptr(mn), ptr(mx),
                                               float(viewingCosLimit), *[ptr(out[k]) for k in
                                               ("mbTrackInView", "mTrackProjX", "mTrackProjY", "mTrackProjXR",
                                                "mnTrackScaleLevel", "mTrackViewCos", "mTrackDepth")]))
    return out


class LocalMap:
    """A local map resident on the device (include/cuorb.h, cuorb_localmap_*): Tracking::SearchLocalPoints' isInFrustum
    loop and ORBmatcher::SearchByProjection(F, mvpLocalMapPoints, ...) in one call per frame."""

    def __init__(self, capacity, device=0):
        self._L = _lib.load()
        self.capacity = int(capacity)
        self._h = self._L.cuorb_localmap_create(int(device), int(capacity))
        if not self._h:
            raise RuntimeError("cuorb_localmap_create: " + _lib.last_error())

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.cuorb_localmap_destroy(self._h)
            self._h = None

    def update(self, first, world_pos, normal, mfMinDistance, mfMaxDistance, descriptors, isBad, hasObservations,
               new_count=None):
        P = np.ascontiguousarray(world_pos, np.float32).reshape(-1, 3); N = np.ascontiguousarray(normal, np.float32).reshape(-1, 3)
        mn = np.ascontiguousarray(mfMinDistance, np.float32); mx = np.ascontiguousarray(mfMaxDistance, np.float32)
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        fl = (np.asarray(isBad, np.uint8) * _lib.MP_BAD | np.asarray(hasObservations, np.uint8) * _lib.MP_HAS_OBS).astype(np.uint8)
        n = len(P)
        cnt = max(self.count, first + n) if new_count is None else int(new_count)
        check(self._L.cuorb_localmap_update(self._h, int(first), n, cnt, ptr(P), ptr(N), ptr(mn), ptr(mx), ptr(d), ptr(fl)))

    @property
    def count(self):
        return check(self._L.cuorb_localmap_count(self._h))

    def search(self, camera, frame, skip_idx=(), viewingCosLimit=0.5, th=1.0, bFarPoints=False, thFarPoints=50.0, nnratio=0.8,
               match_capacity=None, want_best=True, order=None, want_in_view=False):
        """frame: matcher.Frame.  order: storage slots in the order mvpLocalMapPoints holds the points this frame (None =
        storage order); skip_idx and all results are positions in that sequence.  Returns (nmatches, n_in_view,
        match_pos, match_feature, best_idx or None[, in_view]) and replays F.mvpMapPoints[feature] = position in order."""
        import ctypes as C
        fv, keep = frame.view()
        skip = np.ascontiguousarray(np.asarray(skip_idx, np.int32))
        ordr = None if order is None else np.ascontiguousarray(np.asarray(order, np.int32))
        M = self.count if ordr is None else len(ordr)
        cap = int(match_capacity) if match_capacity is not None else max(M, 1)
        mp = np.full(max(cap, 1), -1, np.int32); ft = np.full(max(cap, 1), -1, np.int32)
        best = np.full(max(M, 1), -1, np.int32) if want_best else None
        inv = np.zeros(max(M, 1), np.uint8) if want_in_view else None
        nin = C.c_int32(0)
        n = check(self._L.cuorb_localmap_search(self._h, C.byref(camera), C.byref(fv), 0 if ordr is None else len(ordr),
                                                None if ordr is None else ptr(ordr), len(skip), ptr(skip) if len(skip) else None,
                                                float(viewingCosLimit), float(th), int(bFarPoints), float(thFarPoints),
                                                float(nnratio), ptr(best) if want_best else None,
                                                ptr(inv) if want_in_view else None, cap, ptr(mp), ptr(ft), C.byref(nin)))
        for p, f in zip(mp[:n], ft[:n]):
            frame.mvpMapPoints[f] = p
        out = (n, nin.value, mp[:n].copy(), ft[:n].copy(), (best[:M] if want_best else None))
        return out + (inv[:M],) if want_in_view else out

    def search_extracted(self, camera, frame, extractor, image_index=0, **kw):
        """search() on the Frame the extractor just produced, read where the extractor left it on the device:
        keypoints, descriptors and (when frame.mvuRight is not None, i.e. after SearchStereo on this handle) uRight of
        image `image_index` of extractor's last batch.  `frame` supplies N, the grid constants, the scale factors and
        the bound flags; its mvKeysUn / mDescriptors / mvuRight arrays are NOT read.  Keywords as search_resident."""
        import ctypes as C
        dk, dd, du, dz = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(self._L.cuorb_extractor_device_features(extractor.handle, int(image_index), C.byref(dk), C.byref(dd), C.byref(du), C.byref(dz)))
        if frame.mvuRight is not None and not du.value:
            raise RuntimeError("frame has mvuRight but no stereo match has run on this extractor handle")
        return self.search_resident(camera, frame, dk.value, dd.value, du.value if frame.mvuRight is not None else None,
                                    producer_stream=extractor.stream, **kw)

    def search_resident(self, camera, frame, d_keys, d_descriptors, d_u_right=None, producer_stream=None, K4=None, dist5=None,
                        want_keys_un=False, skip_idx=(), viewingCosLimit=0.5, th=1.0, bFarPoints=False, thFarPoints=50.0,
                        nnratio=0.8, match_capacity=None, want_best=True, order=None, want_in_view=False):
        """cuorb_localmap_search_resident: d_keys / d_descriptors / d_u_right are device addresses (ints) of the Frame's
        mvKeys (or mvKeysUn when K4 / dist5 are None), mDescriptors and mvuRight; producer_stream the cudaStream_t
        (int) that wrote them or None.  K4 / dist5: Frame::UndistortKeyPoints runs on the device first.  Returns
        search()'s tuple, plus mvKeysUn as the search used it when want_keys_un."""
        import ctypes as C
        fv, keep = frame.view()
        fv.keys_un = d_keys; fv.descriptors = d_descriptors; fv.u_right = d_u_right
        k4 = None if K4 is None else np.ascontiguousarray(K4, np.float32)
        d5 = None if dist5 is None else np.ascontiguousarray(dist5, np.float32)
        kun = np.zeros(max(frame.N, 1), _lib.KP_DTYPE) if want_keys_un else None
        skip = np.ascontiguousarray(np.asarray(skip_idx, np.int32))
        ordr = None if order is None else np.ascontiguousarray(np.asarray(order, np.int32))
        M = self.count if ordr is None else len(ordr)
        cap = int(match_capacity) if match_capacity is not None else max(M, 1)
        mp = np.full(max(cap, 1), -1, np.int32); ft = np.full(max(cap, 1), -1, np.int32)
        best = np.full(max(M, 1), -1, np.int32) if want_best else None
        inv = np.zeros(max(M, 1), np.uint8) if want_in_view else None
        nin = C.c_int32(0)
        n = check(self._L.cuorb_localmap_search_resident(
            self._h, C.byref(camera), C.byref(fv), producer_stream or None,
            None if k4 is None else ptr(k4), None if d5 is None else ptr(d5), ptr(kun) if want_keys_un else None,
            0 if ordr is None else len(ordr), None if ordr is None else ptr(ordr), len(skip), ptr(skip) if len(skip) else None,
            float(viewingCosLimit), float(th), int(bFarPoints), float(thFarPoints), float(nnratio),
            ptr(best) if want_best else None, ptr(inv) if want_in_view else None, cap, ptr(mp), ptr(ft), C.byref(nin)))
        for p, f in zip(mp[:n], ft[:n]):
            frame.mvpMapPoints[f] = p
        out = (n, nin.value, mp[:n].copy(), ft[:n].copy(), (best[:M] if want_best else None))
        if want_in_view: out = out + (inv[:M],)
        if want_keys_un: out = out + (kun[:frame.N],)
        return out
