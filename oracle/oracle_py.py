"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

ctypes binding of the CPU oracle (oracle/_build/liborb_oracle.so, built by oracle/Makefile).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module; the product package (orbslam3_cpp_fork_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "liborb_oracle.so")

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
     ("octave", "<i4"), ("class_id", "<i4")]
)
assert KP_DTYPE.itemsize == 28


class FrameViewC(C.Structure):
    """cuorb_frame_view (include/cuorb.h)."""

    _fields_ = [
        ("n", C.c_int32), ("keys_un", C.c_void_p), ("descriptors", C.c_void_p),
        ("u_right", C.c_void_p), ("bound_with_obs", C.c_void_p),
        ("min_x", C.c_float), ("min_y", C.c_float), ("max_x", C.c_float), ("max_y", C.c_float),
        ("grid_w_inv", C.c_float), ("grid_h_inv", C.c_float), ("mb", C.c_float), ("mbf", C.c_float),
        ("nlevels", C.c_int32), ("scale_factors", C.c_void_p),
    ]


class FrustumCamera(C.Structure):
    """cuorb_frustum_camera (include/cuorb.h)."""

    _fields_ = [("rcw", C.c_float * 9), ("tcw", C.c_float * 3), ("ow", C.c_float * 3),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("mbf", C.c_float),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("log_scale_factor", C.c_float), ("n_levels", C.c_int32)]


def make_frustum_camera(Rcw, tcw, K4, mbf, w, h, scale_factor=1.2, n_levels=8):
    """Frame members isInFrustum reads, from a world->camera pose: mOw = -Rcw^T tcw in float32 (row dot products
    associated x0 + (x1 + x2), as everywhere in this statement of the path)."""
    R = np.asarray(Rcw, np.float32).reshape(3, 3); t = np.asarray(tcw, np.float32).reshape(3)
    Rt = R.T
    ow = -np.array([np.float32(Rt[r, 0] * t[0]) + np.float32(np.float32(Rt[r, 1] * t[1]) + np.float32(Rt[r, 2] * t[2]))
                    for r in range(3)], np.float32)
    c = FrustumCamera()
    c.rcw[:] = [float(x) for x in R.ravel()]; c.tcw[:] = [float(x) for x in t]; c.ow[:] = [float(x) for x in ow]
    c.fx, c.fy, c.cx, c.cy = [float(np.float32(x)) for x in K4]
    c.mbf = float(np.float32(mbf))
    c.min_x, c.max_x, c.min_y, c.max_y = 0.0, float(w), 0.0, float(h)
    c.log_scale_factor = float(np.float32(np.log(np.float64(np.float32(scale_factor)))))   # mfLogScaleFactor = log(mfScaleFactor)
    c.n_levels = n_levels
    return c


def _frustum_call(fn, cam, P, N, min_d, max_d, limit):
    P = np.ascontiguousarray(P, np.float32); N = np.ascontiguousarray(N, np.float32)
    min_d = np.ascontiguousarray(min_d, np.float32); max_d = np.ascontiguousarray(max_d, np.float32)
    n = len(min_d)
    out = dict(in_view=np.zeros(n, np.uint8), proj_x=np.zeros(n, np.float32), proj_y=np.zeros(n, np.float32),
               proj_xr=np.zeros(n, np.float32), level=np.zeros(n, np.int32), view_cos=np.zeros(n, np.float32),
               depth=np.zeros(n, np.float32))
    cnt = fn(C.byref(cam), n, _p(P), _p(N), _p(min_d), _p(max_d), C.c_float(limit), *[_p(out[k]) for k in
             ("in_view", "proj_x", "proj_y", "proj_xr", "level", "view_cos", "depth")])
    out["count"] = cnt
    return out


def is_in_frustum(cam, P, N, min_d, max_d, limit):
    return _frustum_call(lib().oo_is_in_frustum, cam, P, N, min_d, max_d, limit)


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in
            ("orb_oracle.cpp", "orb_oracle_match.cpp", "orb_oracle_capi.cpp", "preproc_oracle.cpp", "frameglue_oracle.cpp", "bfmatch_oracle.cpp", "bow_oracle.cpp", "search_oracle.cpp", "orb_oracle.hpp")]
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.oo_create.restype = C.c_void_p
        L.oo_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.oo_destroy.argtypes = [C.c_void_p]
        L.oo_get_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        L.oo_get_umax.argtypes = [C.c_void_p, C.c_void_p]
        L.oo_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int,
                                 C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.oo_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.oo_get_plane.restype = C.c_int64
        L.oo_get_plane.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
        L.oo_get_candidates.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.oo_get_level_keys.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.oo_resize.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_int, C.c_int]
        L.oo_border.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_int]
        L.oo_fast.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.oo_blur.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p]
        L.oo_fast_atan2.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.oo_octree.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.c_int, C.c_void_p, C.c_int]
        L.oo_hamming.argtypes = [C.c_void_p, C.c_void_p]
        L.oo_stereo.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                C.c_void_p, C.c_void_p, C.c_float, C.c_float] + [C.c_void_p] * 5
        L.oo_search_by_projection.argtypes = (
            [C.POINTER(FrameViewC), C.c_int] + [C.c_void_p] * 8 +
            [C.c_float, C.c_int, C.c_float, C.c_float, C.c_void_p])
        L.oo_search_by_projection_frame.argtypes = (
            [C.POINTER(FrameViewC), C.c_int] + [C.c_void_p] * 7 +
            [C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p])
        L.oo_features_in_area.argtypes = [C.POINTER(FrameViewC), C.c_float, C.c_float, C.c_float, C.c_int, C.c_int,
                                          C.c_void_p, C.c_int]
        L.oo_bench_stereo.restype = C.c_double
        L.oo_bench_stereo.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                      C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                                      C.c_int, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.oo_cvt_gray.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
        L.oo_remap.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p,
                               C.c_void_p, C.c_int, C.c_int, C.c_size_t]
        L.oo_resize_cn.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_void_p, C.c_int,
                                   C.c_int, C.c_size_t]
        L.oo_preprocess.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.oo_undistort_keypoints.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.oo_stereo_from_rgbd.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_float,
                                          C.c_void_p, C.c_void_p]
        L.oo_is_in_frustum.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_float] + [C.c_void_p] * 7
        L.oo_knn_match2.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.oo_fisheye_ratio_test.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.oo_distinctive_descriptors.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.oo_search_by_bow.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                       C.c_float, C.c_int, C.c_void_p]
        FV = C.POINTER(FrameViewC)
        L.oo_search_by_projection_reloc.argtypes = [FV, C.c_int] + [C.c_void_p] * 6 + [C.c_float, C.c_int, C.c_int] + [C.c_void_p] * 3
        L.oo_search_by_projection_sim3.argtypes = [FV, C.c_int] + [C.c_void_p] * 5 + [C.c_int, C.c_float, C.c_void_p]
        L.oo_fuse.argtypes = [FV, C.c_void_p, C.c_int] + [C.c_void_p] * 6 + [C.c_float, C.c_int, C.c_void_p]
        L.oo_search_by_sim3.argtypes = [FV, FV] + [C.c_void_p] * 10 + [C.c_float, C.c_void_p]
        L.oo_search_for_initialization.argtypes = [FV, FV, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p]
        L.oo_search_for_triangulation.argtypes = [FV, FV, C.c_int] + [C.c_void_p] * 7 + [C.c_int, C.c_int, C.c_int, C.c_void_p]
        V = C.c_void_p
        L.oo_search_by_projection_rig.argtypes = ([FV, FV, V, V, C.c_int] + [V] * 11 + [C.c_float, C.c_int, C.c_float, C.c_float, V, V])
        L.oo_search_by_projection_frame_rig.argtypes = ([FV, FV, C.c_int] + [V] * 9 + [C.c_float, C.c_int, C.c_int, C.c_int] + [V] * 5)
        L.oo_search_by_bow_rig.argtypes = [C.c_int, V, V, V, C.c_int, C.c_int, V, V, C.c_int, V, V, V, V, C.c_float, C.c_int, V, V]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_frame_view(keys_un, desc, u_right, bound, w, h, scale_factors, mb=0.0, mbf=0.0,
                    min_x=0.0, min_y=0.0, max_x=None, max_y=None):
    """Build a cuorb_frame_view (+ keepalive list) the way the Frame ctor fills the members
    (src/Frame.cc:240-257: mfGridElementWidthInv = 64 / (mnMaxX - mnMinX), ...)."""
    max_x = float(w) if max_x is None else max_x
    max_y = float(h) if max_y is None else max_y
    keys_un = np.ascontiguousarray(keys_un, dtype=KP_DTYPE)
    desc = np.ascontiguousarray(desc, dtype=np.uint8)
    sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
    ur = None if u_right is None else np.ascontiguousarray(u_right, dtype=np.float32)
    bd = None if bound is None else np.ascontiguousarray(bound, dtype=np.uint8)
    fv = FrameViewC()
    fv.n = len(keys_un)
    fv.keys_un = _p(keys_un); fv.descriptors = _p(desc); fv.u_right = _p(ur); fv.bound_with_obs = _p(bd)
    fv.min_x, fv.min_y, fv.max_x, fv.max_y = min_x, min_y, max_x, max_y
    fv.grid_w_inv = float(np.float32(64.0) / np.float32(np.float32(max_x) - np.float32(min_x)))
    fv.grid_h_inv = float(np.float32(48.0) / np.float32(np.float32(max_y) - np.float32(min_y)))
    fv.mb, fv.mbf = mb, mbf
    fv.nlevels = len(sf)
    fv.scale_factors = _p(sf)
    return fv, [keys_un, desc, sf, ur, bd]


class OracleExtractor:
    """ORBextractor of the oracle (same ctor arguments as include/ORBextractor.h:48)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.L = lib()
        self.nlevels = nlevels
        self.h = self.L.oo_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        self.cap = max(4 * nfeatures, 64)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.oo_destroy(self.h)
            self.h = None

    def tables(self):
        n = self.nlevels
        sf, inv, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        per = np.zeros(n, np.int32)
        self.L.oo_get_tables(self.h, _p(sf), _p(inv), _p(s2), _p(is2), _p(per))
        return sf, inv, s2, is2, per

    def umax(self):
        out = np.zeros(16, np.int32)
        self.L.oo_get_umax(self.h, _p(out))
        return out

    def extract(self, img, lap=(0, 0)):
        """operator(): returns (monoIndex, keypoints[KP_DTYPE], descriptors[n,32])."""
        if img is None or img.size == 0:
            return -1, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        img = np.ascontiguousarray(img, dtype=np.uint8)
        h, w = img.shape
        kps = np.zeros(self.cap, KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = C.c_int(0)
        mono = self.L.oo_extract(self.h, _p(img), w, h, img.strides[0], lap[0], lap[1], _p(kps),
                                 _p(desc), self.cap, C.byref(n))
        if mono < 0:
            raise RuntimeError(f"oracle extract failed: {mono}")
        return mono, kps[: n.value].copy(), desc[: n.value].copy()

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        self.L.oo_level_size(self.h, level, C.byref(w), C.byref(h))
        return w.value, h.value

    def plane(self, level, plane):
        w, h = self.level_size(level)
        if plane == 0:
            out = np.zeros((h + 38, w + 38), np.uint8)
        else:
            out = np.zeros((h, w), np.uint8)
        r = self.L.oo_get_plane(self.h, level, plane, _p(out), out.size)
        if r == 0:
            return None
        assert r == out.size, r
        return out

    def candidates(self, level, cap=200000):
        xs, ys, rs = (np.zeros(cap, np.int32) for _ in range(3))
        n = self.L.oo_get_candidates(self.h, level, _p(xs), _p(ys), _p(rs), cap)
        assert n <= cap
        return xs[:n].copy(), ys[:n].copy(), rs[:n].copy()

    def level_keys(self, level):
        out = np.zeros(self.cap, KP_DTYPE)
        n = self.L.oo_get_level_keys(self.h, level, _p(out), self.cap)
        return out[:n].copy()

    def octree(self, keys, min_x, max_x, min_y, max_y, n_desired):
        keys = np.ascontiguousarray(keys, dtype=KP_DTYPE)
        cap = max(len(keys), 16)
        out = np.zeros(cap, KP_DTYPE)
        n = self.L.oo_octree(self.h, _p(keys), len(keys), min_x, max_x, min_y, max_y, n_desired,
                             _p(out), cap)
        return out[:n].copy()


def resize(src, dw, dh):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros((dh, dw), np.uint8)
    lib().oo_resize(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), dw, dh)
    return dst


def border(src, b=19):
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros((src.shape[0] + 2 * b, src.shape[1] + 2 * b), np.uint8)
    lib().oo_border(_p(src), src.shape[1], src.shape[0], src.strides[0], _p(dst), b)
    return dst


def fast(img, threshold, nonmax=True, cap=400000):
    img = np.ascontiguousarray(img, np.uint8)
    out = np.zeros(cap, KP_DTYPE)
    n = lib().oo_fast(_p(img), img.shape[1], img.shape[0], img.strides[0], threshold, int(nonmax),
                      _p(out), cap)
    assert n <= cap
    return out[:n].copy()


def blur(img):
    img = np.ascontiguousarray(img, np.uint8)
    dst = np.zeros_like(img)
    lib().oo_blur(_p(img), img.shape[1], img.shape[0], img.strides[0], _p(dst))
    return dst


def fast_atan2(y, x):
    y = np.ascontiguousarray(y, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros_like(y)
    lib().oo_fast_atan2(_p(y), _p(x), _p(out), y.size)
    return out


# ---- pre-extraction image ops (preproc_oracle.cpp) ----
PIX_GRAY, PIX_RGB, PIX_BGR, PIX_RGBA, PIX_BGRA = range(5)
PRE_NONE, PRE_REMAP, PRE_RESIZE = range(3)


def _channels(a):
    return 1 if a.ndim == 2 else a.shape[2]


def cvt_gray(img, rgb_order):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w, ch = img.shape
    out = np.empty((h, w), np.uint8)
    lib().oo_cvt_gray(_p(img), w, h, w * ch, ch, int(bool(rgb_order)), _p(out), w)
    return out


def remap(img, mapx, mapy):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    mapx = np.ascontiguousarray(mapx, dtype=np.float32); mapy = np.ascontiguousarray(mapy, dtype=np.float32)
    ch = _channels(img)
    dh, dw = mapx.shape
    out = np.empty((dh, dw) if ch == 1 else (dh, dw, ch), np.uint8)
    lib().oo_remap(_p(img), img.shape[1], img.shape[0], img.shape[1] * ch, ch, _p(mapx), _p(mapy), _p(out), dw, dh, dw * ch)
    return out


def resize_cn(img, dw, dh):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    ch = _channels(img)
    out = np.empty((dh, dw) if ch == 1 else (dh, dw, ch), np.uint8)
    lib().oo_resize_cn(_p(img), img.shape[1], img.shape[0], img.shape[1] * ch, ch, _p(out), dw, dh, dw * ch)
    return out


def preprocess(img, fmt, mode=PRE_NONE, mapx=None, mapy=None, dsize=None):
    """remap|resize (on the raw channels) then gray conversion, as System::Track* + Tracking::GrabImage* do."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    ch = _channels(img)
    sh, sw = img.shape[:2]
    if mode == PRE_REMAP:
        mapx = np.ascontiguousarray(mapx, dtype=np.float32); mapy = np.ascontiguousarray(mapy, dtype=np.float32)
        dh, dw = mapx.shape
    elif mode == PRE_RESIZE:
        dw, dh = dsize
    else:
        dw, dh = sw, sh
    out = np.empty((dh, dw), np.uint8)
    lib().oo_preprocess(_p(img), sw, sh, sw * ch, int(fmt), int(mode), _p(mapx), _p(mapy), _p(out), dw, dh)
    return out


# ---- post-extraction frame glue (frameglue_oracle.cpp) ----
def undistort_keypoints(kps, K4, dist):
    """Frame::UndistortKeyPoints: K4 = (fx, fy, cx, cy), dist = (k1, k2, p1, p2[, k3])."""
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
    out = np.empty_like(kps)
    K4 = np.ascontiguousarray(K4, dtype=np.float32)
    d5 = np.zeros(5, np.float32); d5[:len(dist)] = np.asarray(dist, np.float32)
    lib().oo_undistort_keypoints(_p(kps), _p(out), len(kps), _p(K4), _p(d5))
    return out


def stereo_from_rgbd(kps, kps_un, depth, mbf):
    """Frame::ComputeStereoFromRGBD: depth float32 [h, w] -> (mvuRight, mvDepth)."""
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE); kps_un = np.ascontiguousarray(kps_un, dtype=KP_DTYPE)
    depth = np.ascontiguousarray(depth, dtype=np.float32)
    u = np.empty(len(kps), np.float32); d = np.empty(len(kps), np.float32)
    lib().oo_stereo_from_rgbd(_p(kps), _p(kps_un), len(kps), _p(depth), depth.shape[1], np.float32(mbf), _p(u), _p(d))
    return u, d


# ---- dense all-pairs Hamming (bfmatch_oracle.cpp) ----
def knn_match2(q, t):
    """cv::BFMatcher(NORM_HAMMING).knnMatch(q, t, k=2) -> (idx [nq, 2], dist [nq, 2]); -1 where missing."""
    q = np.ascontiguousarray(q, dtype=np.uint8).reshape(-1, 32); t = np.ascontiguousarray(t, dtype=np.uint8).reshape(-1, 32)
    idx = np.empty((len(q), 2), np.int32); dist = np.empty((len(q), 2), np.int32)
    lib().oo_knn_match2(_p(q), len(q), _p(t), len(t), _p(idx), _p(dist))
    return idx, dist


def fisheye_ratio_test(idx, dist):
    idx = np.ascontiguousarray(idx, dtype=np.int32); dist = np.ascontiguousarray(dist, dtype=np.int32)
    out = np.empty(len(idx), np.int32)
    lib().oo_fisheye_ratio_test(_p(idx), _p(dist), len(idx), _p(out))
    return out


def distinctive_descriptors(desc, off):
    """MapPoint::ComputeDistinctiveDescriptors for many points: desc [sum n, 32], off [n_points + 1]."""
    desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32); off = np.ascontiguousarray(off, dtype=np.int32)
    best = np.empty(len(off) - 1, np.int32)
    lib().oo_distinctive_descriptors(_p(desc), _p(off), len(off) - 1, _p(best))
    return best


# ---- vocabulary-node searches (bow_oracle.cpp) ----
def search_by_bow(desc1, valid1, angle1, desc2, eligible2, angle2, off1, idx1, off2, idx2, mode, nnratio=0.7,
                  check_orientation=True):
    """ORBmatcher::SearchByBoW on flattened inputs; mode 0 = (KeyFrame, Frame), 1 = (KeyFrame, KeyFrame).
    Returns (nmatches, match12 [n1])."""
    desc1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32); desc2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
    valid1 = np.ascontiguousarray(valid1, np.uint8); eligible2 = np.ascontiguousarray(eligible2, np.uint8)
    angle1 = np.ascontiguousarray(angle1, np.float32); angle2 = np.ascontiguousarray(angle2, np.float32)
    off1 = np.ascontiguousarray(off1, np.int32); idx1 = np.ascontiguousarray(idx1, np.int32)
    off2 = np.ascontiguousarray(off2, np.int32); idx2 = np.ascontiguousarray(idx2, np.int32)
    m = np.empty(len(desc1), np.int32)
    n = lib().oo_search_by_bow(len(desc1), _p(desc1), _p(valid1), _p(angle1), len(desc2), _p(desc2), _p(eligible2), _p(angle2),
                               len(off1) - 1, _p(off1), _p(idx1), _p(off2), _p(idx2), int(mode), np.float32(nnratio),
                               int(bool(check_orientation)), _p(m))
    return n, m


def hamming(a, b):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().oo_hamming(_p(a), _p(b))


def stereo(ext_l: OracleExtractor, ext_r: OracleExtractor, kps_l, desc_l, kps_r, desc_r, mb, mbf):
    """Frame::ComputeStereoMatches on the pyramids of the two extractors' last extract.
    Returns dict(u_right, depth, best_idx_r, best_dist, sad)."""
    kps_l = np.ascontiguousarray(kps_l, KP_DTYPE); kps_r = np.ascontiguousarray(kps_r, KP_DTYPE)
    desc_l = np.ascontiguousarray(desc_l, np.uint8); desc_r = np.ascontiguousarray(desc_r, np.uint8)
    n = len(kps_l)
    u = np.zeros(n, np.float32); d = np.zeros(n, np.float32)
    bi, bd, sad = (np.zeros(n, np.int32) for _ in range(3))
    lib().oo_stereo(ext_l.h, ext_r.h, n, _p(kps_l), _p(desc_l), len(kps_r), _p(kps_r), _p(desc_r),
                    mb, mbf, _p(u), _p(d), _p(bi), _p(bd), _p(sad))
    return dict(u_right=u, depth=d, best_idx_r=bi, best_dist=bd, sad=sad)


def search_by_projection(fv, mp_desc, px, py, pxr, level, view_cos, depth, flags, th, far_points,
                         th_far, nn_ratio):
    M = len(px)
    arrs = [np.ascontiguousarray(mp_desc, np.uint8), np.ascontiguousarray(px, np.float32),
            np.ascontiguousarray(py, np.float32), np.ascontiguousarray(pxr, np.float32),
            np.ascontiguousarray(level, np.int32), np.ascontiguousarray(view_cos, np.float32),
            np.ascontiguousarray(depth, np.float32), np.ascontiguousarray(flags, np.uint8)]
    out = np.zeros(M, np.int32)
    n = lib().oo_search_by_projection(C.byref(fv), M, *[_p(a) for a in arrs], th, int(far_points),
                                      th_far, nn_ratio, _p(out))
    return n, out


def search_by_projection_frame(fv, last_desc, u, v, inv_z, last_octave, last_angle, flags, th,
                               forward, backward, check_ori=True):
    n_last = len(u)
    arrs = [np.ascontiguousarray(last_desc, np.uint8), np.ascontiguousarray(u, np.float32),
            np.ascontiguousarray(v, np.float32), np.ascontiguousarray(inv_z, np.float32),
            np.ascontiguousarray(last_octave, np.int32), np.ascontiguousarray(last_angle, np.float32),
            np.ascontiguousarray(flags, np.uint8)]
    best = np.zeros(n_last, np.int32); rot = np.zeros(n_last, np.int32); keep = np.zeros(3, np.int32)
    n = lib().oo_search_by_projection_frame(C.byref(fv), n_last, *[_p(a) for a in arrs], th,
                                            int(forward), int(backward), int(check_ori), _p(best),
                                            _p(rot), _p(keep))
    return n, best, rot, keep


def bench_stereo(left, right, mb, mbf, n_threads, nfeatures=1000, scale=1.2, nlevels=8, ini=20,
                 mn=7):
    """Time the oracle's stereo path over a stack of frames [n,h,w]; returns (seconds, kps, matches)."""
    left = np.ascontiguousarray(left, np.uint8); right = np.ascontiguousarray(right, np.uint8)
    n, h, w = left.shape
    k, m = C.c_int64(0), C.c_int64(0)
    sec = lib().oo_bench_stereo(nfeatures, scale, nlevels, ini, mn, _p(left), _p(right), n, w, h, mb,
                                mbf, n_threads, C.byref(k), C.byref(m))
    return sec, k.value, m.value


# ---- the remaining ORBmatcher searches (search_oracle.cpp) ----
def _f32(a): return None if a is None else np.ascontiguousarray(a, np.float32)
def _i32(a): return None if a is None else np.ascontiguousarray(a, np.int32)
def _u8(a): return None if a is None else np.ascontiguousarray(a, np.uint8)


def search_by_projection_reloc(fv, mp_desc, u, v, level, kf_angle, active, th, orb_dist, check_ori=True):
    n = len(u)
    arrs = [_u8(mp_desc), _f32(u), _f32(v), _i32(level), _f32(kf_angle), _u8(active)]
    best = np.empty(n, np.int32); rot = np.empty(n, np.int32); keep = np.empty(3, np.int32)
    r = lib().oo_search_by_projection_reloc(C.byref(fv), n, *[_p(a) for a in arrs], th, int(orb_dist), int(check_ori),
                                            _p(best), _p(rot), _p(keep))
    return r, best, rot, keep


def search_by_projection_sim3(fv, mp_desc, u, v, level, active, th, ratio_hamming):
    n = len(u)
    arrs = [_u8(mp_desc), _f32(u), _f32(v), _i32(level), _u8(active)]
    best = np.empty(n, np.int32)
    r = lib().oo_search_by_projection_sim3(C.byref(fv), n, *[_p(a) for a in arrs], int(th), ratio_hamming, _p(best))
    return r, best


def fuse(fv, inv_sigma2, mp_desc, u, v, ur, level, active, th, check_reprojection):
    n = len(u)
    arrs = [_u8(mp_desc), _f32(u), _f32(v), _f32(ur), _i32(level), _u8(active)]
    best = np.empty(n, np.int32)
    r = lib().oo_fuse(C.byref(fv), _p(_f32(inv_sigma2)), n, *[_p(a) for a in arrs], th, int(check_reprojection), _p(best))
    return r, best


def search_by_sim3(fv1, fv2, d1, u12, v12, l12, a1, d2, u21, v21, l21, a2, th):
    arrs = [_u8(d1), _f32(u12), _f32(v12), _i32(l12), _u8(a1), _u8(d2), _f32(u21), _f32(v21), _i32(l21), _u8(a2)]
    m12 = np.empty(max(fv1.n, 1), np.int32)
    r = lib().oo_search_by_sim3(C.byref(fv1), C.byref(fv2), *[_p(a) for a in arrs], th, _p(m12))
    return r, m12[:fv1.n]


def search_for_initialization(fv1, fv2, prev_matched, window, nnratio, check_ori=True):
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.empty(max(fv1.n, 1), np.int32)
    r = lib().oo_search_for_initialization(C.byref(fv1), C.byref(fv2), _p(pm), int(window), nnratio, int(check_ori), _p(m12))
    return r, m12[:fv1.n], pm


def search_for_triangulation(fv1, fv2, off1, idx1, off2, idx2, F12, ep, sigma2_2, only_stereo, coarse, check_ori=True):
    arrs = [_i32(off1), _i32(idx1), _i32(off2), _i32(idx2), _f32(F12), _f32(ep), _f32(sigma2_2)]
    m12 = np.empty(max(fv1.n, 1), np.int32)
    r = lib().oo_search_for_triangulation(C.byref(fv1), C.byref(fv2), len(off1) - 1, *[_p(a) for a in arrs],
                                          int(only_stereo), int(coarse), int(check_ori), _p(m12))
    return r, m12[:fv1.n]


# ---- two-camera (Frame::Nleft != -1) branches of the tracking-time searches (rig_oracle.cpp) ----
MP_IN_VIEW_R = 8


def search_by_projection_rig(fvl, fvr, l2r, r2l, mp_desc, px, py, level, view_cos, pxr, pyr, level_r, view_cos_r, depth, flags,
                             th, far_points, th_far, nn_ratio):
    """Returns (nmatches incl. the cross-bound stereo partners, best_l[M], best_r[M])."""
    arrs = [_u8(mp_desc), _f32(px), _f32(py), _i32(level), _f32(view_cos), _f32(pxr), _f32(pyr), _i32(level_r),
            _f32(view_cos_r), _f32(depth), _u8(flags)]
    a, b = _i32(l2r), _i32(r2l)
    M = len(arrs[1])
    bl = np.empty(max(M, 1), np.int32); br = np.empty(max(M, 1), np.int32)
    n = lib().oo_search_by_projection_rig(C.byref(fvl), C.byref(fvr), _p(a), _p(b), M, *[_p(x) for x in arrs], th, int(far_points),
                                          th_far, nn_ratio, _p(bl), _p(br))
    return n, bl[:M], br[:M]


def search_by_projection_frame_rig(fvl, fvr, last_desc, u, v, inv_z, ur, vr, last_octave, last_angle, flags, th, forward, backward,
                                   check_ori=True):
    """Returns (nmatches, best_l, best_r, rot_bin_l, rot_bin_r, keep_bins)."""
    arrs = [_u8(last_desc), _f32(u), _f32(v), _f32(inv_z), _f32(ur), _f32(vr), _i32(last_octave), _f32(last_angle), _u8(flags)]
    n_last = len(arrs[1])
    outs = [np.empty(max(n_last, 1), np.int32) for _ in range(4)] + [np.empty(3, np.int32)]
    n = lib().oo_search_by_projection_frame_rig(C.byref(fvl), C.byref(fvr), n_last, *[_p(x) for x in arrs], th, int(forward),
                                                int(backward), int(check_ori), *[_p(o) for o in outs])
    return (n,) + tuple(o[:n_last] for o in outs[:4]) + (outs[4],)


def search_by_bow_rig(desc1, valid1, angle1, desc2, nleft2, angle2, off1, idx1, off2, idx2, nnratio=0.7, check_orientation=True):
    """SearchByBoW(KeyFrame, Frame) with F.Nleft != -1; side-2 arrays cover F's left-then-right features.
    Returns (nmatches, match_l[n1], match_r[n1])."""
    desc1 = _u8(desc1).reshape(-1, 32); desc2 = _u8(desc2).reshape(-1, 32)
    v1 = _u8(np.asarray(valid1).astype(np.uint8)); a1 = _f32(angle1); a2 = _f32(angle2)
    o1, i1, o2, i2 = _i32(off1), _i32(idx1), _i32(off2), _i32(idx2)
    ml = np.empty(max(len(desc1), 1), np.int32); mr = np.empty(max(len(desc1), 1), np.int32)
    n = lib().oo_search_by_bow_rig(len(desc1), _p(desc1), _p(v1), _p(a1), len(desc2), int(nleft2), _p(desc2), _p(a2), len(o1) - 1,
                                   _p(o1), _p(i1), _p(o2), _p(i2), np.float32(nnratio), int(bool(check_orientation)), _p(ml), _p(mr))
    return n, ml[:len(desc1)], mr[:len(desc1)]
