"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

ctypes binding of oracle/_ref/liborb_ref.so: the reference's OWN sources (unmodified
src/ORBextractor.cc and src/ORBmatcher.cc, plus function bodies lifted verbatim out of src/Frame.cc,
src/KeyFrame.cc, src/MapPoint.cc, src/CameraModels/Pinhole.cpp) compiled over the stand-in headers of
oracle/ref_compat (recipe: `make -C oracle ref`).  It is what pins the restated oracle
(tests/test_oracle_vs_ref.py) and, through it or directly, libcuorb (tests/test_gpu_ref_parity.py).

/root/reference only exists in the build container: there the library is (re)built on demand; on the
GPU box the prebuilt, git-ignored oracle/_ref/liborb_ref.so that travelled with the snapshot is
loaded as it is.  Only tests/, __graft_entry__ and bench.py's CPU legs import this module.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess

import numpy as np

from .oracle_py import KP_DTYPE, FrameViewC, _p

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_ref", "liborb_ref.so")


def reference_root() -> str:
    try:
        with open(os.path.join(_HERE, "..", "BASELINE.json")) as f:
            return json.load(f).get("reference_path", "/root/reference")
    except OSError:
        return "/root/reference"


def available() -> bool:
    return os.path.exists(_LIB_PATH) or os.path.isdir(os.path.join(reference_root(), "src"))


def build(force: bool = False) -> str | None:
    """Build oracle/_ref when the reference checkout is present; otherwise keep the prebuilt file."""
    ref = reference_root()
    if os.path.isfile(os.path.join(ref, "src", "ORBextractor.cc")):
        subprocess.check_call(["make", "-C", _HERE, "-s", "ref", f"REF={ref}"] + (["-B"] if force else []))
    return _LIB_PATH if os.path.exists(_LIB_PATH) else None


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if build() is None:
            raise RuntimeError("oracle/_ref/liborb_ref.so is missing and the reference checkout is not here to build it")
        L = C.CDLL(_LIB_PATH)
        FV = C.POINTER(FrameViewC)
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_destroy.argtypes = [C.c_void_p]
        L.ref_get_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        L.ref_get_umax.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.ref_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.ref_get_plane.restype = C.c_int64
        L.ref_get_plane.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
        L.ref_get_level_keys.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.ref_octree.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                 C.c_void_p, C.c_int]
        L.ref_hamming.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_stereo.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                 C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.ref_features_in_area.argtypes = [FV, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.ref_grid_cell.argtypes = [FV, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.ref_search_by_projection.argtypes = ([FV, C.c_int] + [C.c_void_p] * 8 +
                                               [C.c_float, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p])
        L.ref_search_by_projection_frame.argtypes = ([FV, C.c_int] + [C.c_void_p] * 7 +
                                                     [C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p])
        L.ref_search_by_projection_frame_pose.argtypes = ([FV] + [C.c_void_p] * 3 + [C.c_int] + [C.c_void_p] * 7 +
                                                          [C.c_float, C.c_int, C.c_int, C.c_void_p])
        L.ref_is_in_frustum.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_float] + [C.c_void_p] * 7
        L.ref_bench_stereo.restype = C.c_double
        L.ref_bench_stereo.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_int,
                                       C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        V = C.c_void_p
        L.ref_search_by_bow.argtypes = ([C.c_int] + [FV, V, C.c_int, V, V, V] * 2 + [C.c_float, C.c_int, V])
        L.ref_search_for_initialization.argtypes = [FV, FV, V, C.c_int, C.c_float, C.c_int, V]
        L.ref_search_by_projection_reloc.argtypes = [FV, C.c_int] + [V] * 7 + [C.c_float, C.c_int, C.c_int, V]
        L.ref_search_by_projection_sim3.argtypes = [FV, C.c_int] + [V] * 5 + [C.c_int, C.c_float, C.c_int, V]
        L.ref_fuse.argtypes = [FV, C.c_int] + [V] * 6 + [C.c_float, C.c_int, V, V, V, V]
        L.ref_search_by_sim3.argtypes = [FV, FV] + [V] * 10 + [C.c_float, V, V]
        L.ref_search_for_triangulation.argtypes = ([FV, V, C.c_int, V, V, V] * 2 + [V, C.c_int, C.c_int, C.c_int, V, V, V])
        L.ref_search_by_projection_rig.argtypes = ([FV, FV, V, V, C.c_int] + [V] * 11 + [C.c_float, C.c_int, C.c_float, C.c_float, V, V])
        L.ref_search_by_projection_frame_rig.argtypes = ([FV, FV, C.c_int] + [V] * 4 + [C.c_float, C.c_float] + [V] * 3 +
                                                         [C.c_float, C.c_int, C.c_int, C.c_int, V])
        L.ref_fisheye_stereo.argtypes = [C.c_int, V, V, C.c_int, C.c_int, V, V, C.c_int, C.c_int, V, V, V, V, V]
        L.ref_search_by_bow_rig.argtypes = [FV, FV, V, C.c_int, V, V, V, FV, FV, C.c_int, V, V, V, C.c_float, C.c_int, V, V]
        L.ref_fuse_rig.argtypes = [FV, FV, C.c_int, C.c_int] + [V] * 6 + [C.c_float, V]
        _lib = L
    return _lib


class RefExtractor:
    """The reference's ORBextractor (same ctor arguments as include/ORBextractor.h:48)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.L = lib()
        self.nlevels = nlevels
        self.h = self.L.ref_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        self.cap = max(4 * nfeatures, 64)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_destroy(self.h)
            self.h = None

    def tables(self):
        n = self.nlevels
        sf, inv, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        per = np.zeros(n, np.int32)
        self.L.ref_get_tables(self.h, _p(sf), _p(inv), _p(s2), _p(is2), _p(per))
        return sf, inv, s2, is2, per

    def umax(self):
        out = np.zeros(16, np.int32)
        self.L.ref_get_umax(self.h, _p(out))
        return out

    def extract(self, img, lap=(0, 0)):
        """operator(): returns (monoIndex, keypoints[KP_DTYPE], descriptors[n,32])."""
        if img is None or img.size == 0:
            n = C.c_int(0)
            r = self.L.ref_extract(self.h, None, 0, 0, 0, lap[0], lap[1], None, None, 0, C.byref(n))
            return r, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        img = np.ascontiguousarray(img, dtype=np.uint8)
        h, w = img.shape
        kps = np.zeros(self.cap, KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = C.c_int(0)
        mono = self.L.ref_extract(self.h, _p(img), w, h, img.strides[0], lap[0], lap[1], _p(kps), _p(desc),
                                  self.cap, C.byref(n))
        if mono < 0:
            raise RuntimeError(f"reference extract failed: {mono}")
        return mono, kps[: n.value].copy(), desc[: n.value].copy()

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        self.L.ref_level_size(self.h, level, C.byref(w), C.byref(h))
        return w.value, h.value

    def plane(self, level, plane):
        """0: bordered level (temp), 1: level ROI.  The blurred copy is a local of operator()."""
        w, h = self.level_size(level)
        out = np.zeros((h + 38, w + 38) if plane == 0 else (h, w), np.uint8)
        r = self.L.ref_get_plane(self.h, level, plane, _p(out), out.size)
        assert r == out.size, r
        return out

    def level_keys(self, level):
        out = np.zeros(self.cap, KP_DTYPE)
        n = self.L.ref_get_level_keys(self.h, level, _p(out), self.cap)
        return out[:n].copy()

    def octree(self, keys, min_x, max_x, min_y, max_y, n_desired):
        keys = np.ascontiguousarray(keys, dtype=KP_DTYPE)
        cap = max(len(keys), 16)
        out = np.zeros(cap, KP_DTYPE)
        n = self.L.ref_octree(self.h, _p(keys), len(keys), min_x, max_x, min_y, max_y, n_desired, _p(out), cap)
        return out[:n].copy()


def hamming(a, b):
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    return lib().ref_hamming(_p(a), _p(b))


def stereo(ext_l: RefExtractor, ext_r: RefExtractor, kps_l, desc_l, kps_r, desc_r, mb, mbf):
    """Frame::ComputeStereoMatches (src/Frame.cc:944-1113, the reference's own body) on the pyramids of the
    two extractors' last extract.  Returns (mvuRight, mvDepth)."""
    kps_l = np.ascontiguousarray(kps_l, KP_DTYPE); kps_r = np.ascontiguousarray(kps_r, KP_DTYPE)
    desc_l = np.ascontiguousarray(desc_l, np.uint8); desc_r = np.ascontiguousarray(desc_r, np.uint8)
    n = len(kps_l)
    u = np.full(n, -1, np.float32); d = np.full(n, -1, np.float32)
    lib().ref_stereo(ext_l.h, ext_r.h, n, _p(kps_l), _p(desc_l), len(kps_r), _p(kps_r), _p(desc_r), mb, mbf,
                     _p(u), _p(d))
    return u, d


def features_in_area(fv, x, y, r, min_level=-1, max_level=-1, cap=4096):
    out = np.zeros(cap, np.int32)
    n = lib().ref_features_in_area(C.byref(fv), x, y, r, min_level, max_level, _p(out), cap)
    return out[:n].copy()


def grid_cell(fv, ix, iy, cap=4096):
    out = np.zeros(cap, np.int32)
    n = lib().ref_grid_cell(C.byref(fv), ix, iy, _p(out), cap)
    return out[:n].copy()


def search_by_projection(fv, mp_desc, px, py, pxr, level, view_cos, depth, flags, th, far_points, th_far,
                         nn_ratio):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints).
    Returns (nmatches, best_idx[M], final_binding[N])."""
    M = len(px)
    arrs = [np.ascontiguousarray(mp_desc, np.uint8), np.ascontiguousarray(px, np.float32),
            np.ascontiguousarray(py, np.float32), np.ascontiguousarray(pxr, np.float32),
            np.ascontiguousarray(level, np.int32), np.ascontiguousarray(view_cos, np.float32),
            np.ascontiguousarray(depth, np.float32), np.ascontiguousarray(flags, np.uint8)]
    best = np.zeros(M, np.int32); final = np.zeros(fv.n, np.int32)
    n = lib().ref_search_by_projection(C.byref(fv), M, *[_p(a) for a in arrs], th, int(far_points), th_far,
                                       nn_ratio, _p(best), _p(final))
    return n, best, final


def search_by_projection_frame(fv, last_desc, u, v, z, last_octave, last_angle, flags, th, forward, backward,
                               check_ori=True):
    """ORBmatcher::SearchByProjection(Cur, Last, th, bMono=false).  z = camera-frame depth (the reference
    computes 1.0 / z itself).  Returns (nmatches, final_binding[N]: last-frame feature bound to each current
    feature after the rotation-consistency pass, -1 none)."""
    n_last = len(u)
    arrs = [np.ascontiguousarray(last_desc, np.uint8), np.ascontiguousarray(u, np.float32),
            np.ascontiguousarray(v, np.float32), np.ascontiguousarray(z, np.float32),
            np.ascontiguousarray(last_octave, np.int32), np.ascontiguousarray(last_angle, np.float32),
            np.ascontiguousarray(flags, np.uint8)]
    final = np.zeros(fv.n, np.int32)
    n = lib().ref_search_by_projection_frame(C.byref(fv), n_last, *[_p(a) for a in arrs], th, int(forward),
                                             int(backward), int(check_ori), _p(final))
    return n, final


def bench_stereo(left, right, mb, mbf, n_threads, nfeatures=1000, scale=1.2, nlevels=8, ini=20, mn=7):
    """Time the reference's stereo path over a stack of frames [n,h,w]; returns (seconds, kps, matches)."""
    left = np.ascontiguousarray(left, np.uint8); right = np.ascontiguousarray(right, np.uint8)
    n, h, w = left.shape
    k, m = C.c_int64(0), C.c_int64(0)
    sec = lib().ref_bench_stereo(nfeatures, scale, nlevels, ini, mn, _p(left), _p(right), n, w, h, mb, mbf,
                                 n_threads, C.byref(k), C.byref(m))
    return sec, k.value, m.value


def is_in_frustum(cam, P, N, min_d, max_d, limit):
    """Frame::isInFrustum for n MapPoints; cam: a ctypes cuorb_frustum_camera (oracle_py.FrustumCamera)."""
    from .oracle_py import _frustum_call
    return _frustum_call(lib().ref_is_in_frustum, cam, P, N, min_d, max_d, limit)


def search_by_projection_frame_pose(fv, pose_cur, pose_last, K4, last_desc, world_pos, last_octave, last_angle, has_mp,
                                    outlier, n_obs, th, mono, check_ori=True):
    """ORBmatcher::SearchByProjection(Cur, Last, th, bMono) with its prologue live (poses = 12 floats: R row-major, t).
    Returns (nmatches, final_binding[N])."""
    arrs = [np.ascontiguousarray(pose_cur, np.float32), np.ascontiguousarray(pose_last, np.float32),
            np.ascontiguousarray(K4, np.float32)]
    per = [np.ascontiguousarray(last_desc, np.uint8), np.ascontiguousarray(world_pos, np.float32),
           np.ascontiguousarray(last_octave, np.int32), np.ascontiguousarray(last_angle, np.float32),
           np.ascontiguousarray(has_mp, np.uint8), np.ascontiguousarray(outlier, np.uint8),
           np.ascontiguousarray(n_obs, np.int32)]
    final = np.zeros(fv.n, np.int32)
    n = lib().ref_search_by_projection_frame_pose(C.byref(fv), *[_p(a) for a in arrs], len(last_octave),
                                                  *[_p(a) for a in per], th, int(mono), int(check_ori), _p(final))
    return n, final


# ---- the ORBmatcher searches outside the tracking path (ref_compat/ref_capi_search.cpp) ----
def _u8(a): return None if a is None else np.ascontiguousarray(a, np.uint8)
def _f32(a): return None if a is None else np.ascontiguousarray(a, np.float32)
def _i32(a): return None if a is None else np.ascontiguousarray(a, np.int32)


def flatten_feature_vector(fv: dict):
    """{node: [feature, ...]} -> (node ids ascending, offsets, features in insertion order), int32."""
    nodes = sorted(fv)
    off = np.zeros(len(nodes) + 1, np.int32)
    for k, nd in enumerate(nodes):
        off[k + 1] = off[k] + len(fv[nd])
    idx = np.asarray([i for nd in nodes for i in fv[nd]], np.int32)
    return len(nodes), np.asarray(nodes, np.int32), off, idx


def search_by_bow(mode, fv1, has_mp1, featvec1, fv2, has_mp2, featvec2, nnratio, check_ori=True):
    """ORBmatcher::SearchByBoW, mode 0 = (KeyFrame, Frame), 1 = (KeyFrame, KeyFrame); has_mp: 0 none, 1 good, 2 bad.
    Returns (nmatches, match12[n1])."""
    n1, nd1, o1, i1 = flatten_feature_vector(featvec1)
    n2, nd2, o2, i2 = flatten_feature_vector(featvec2)
    h1 = _u8(has_mp1); h2 = _u8(has_mp2)
    m = np.empty(max(fv1.n, 1), np.int32)
    r = lib().ref_search_by_bow(int(mode), C.byref(fv1), _p(h1), n1, _p(nd1), _p(o1), _p(i1), C.byref(fv2),
                                _p(h2) if h2 is not None else None, n2, _p(nd2), _p(o2), _p(i2), nnratio,
                                int(check_ori), _p(m))
    return r, m[:fv1.n]


def search_for_initialization(fv1, fv2, prev_matched, window, nnratio, check_ori=True):
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m = np.empty(max(fv1.n, 1), np.int32)
    r = lib().ref_search_for_initialization(C.byref(fv1), C.byref(fv2), _p(pm), int(window), nnratio, int(check_ori), _p(m))
    return r, m[:fv1.n], pm


def search_by_projection_reloc(fv, mp_desc, u, v, level, kf_angle, active, found, th, orb_dist, check_ori=True):
    """Returns (nmatches, binding[N]): binding[i2] = query bound to current feature i2 (-1 none, -2 pre-existing)."""
    arrs = [_u8(mp_desc), _f32(u), _f32(v), _i32(level), _f32(kf_angle), _u8(active), _u8(found)]
    out = np.empty(max(fv.n, 1), np.int32)
    r = lib().ref_search_by_projection_reloc(C.byref(fv), len(arrs[1]), *[None if a is None else _p(a) for a in arrs],
                                             th, int(orb_dist), int(check_ori), _p(out))
    return r, out[:fv.n]


def search_by_projection_sim3(fv, mp_desc, u, v, level, active, th, ratio_hamming, with_kfs=False):
    """Returns (nmatches, matched[N]): matched[idx] = query matched to KeyFrame feature idx (-1 none, -2 pre-existing)."""
    arrs = [_u8(mp_desc), _f32(u), _f32(v), _i32(level), _u8(active)]
    out = np.empty(max(fv.n, 1), np.int32)
    r = lib().ref_search_by_projection_sim3(C.byref(fv), len(arrs[1]), *[_p(a) for a in arrs], int(th), ratio_hamming,
                                            int(with_kfs), _p(out))
    return r, out[:fv.n]


def fuse(fv, mp_desc, u, v, z, level, active, th, sim3, n_obs_kf=None, n_obs_q=None):
    """Returns (nFused of one call over all queries, best_idx[n], outcome[n])."""
    arrs = [_u8(mp_desc), _f32(u), _f32(v), _f32(z), _i32(level), _u8(active)]
    n = len(arrs[1])
    best = np.empty(max(n, 1), np.int32); outc = np.empty(max(n, 1), np.int32)
    ok, oq = _i32(n_obs_kf), _i32(n_obs_q)
    r = lib().ref_fuse(C.byref(fv), n, *[_p(a) for a in arrs], th, int(sim3), None if ok is None else _p(ok),
                       None if oq is None else _p(oq), _p(best), _p(outc))
    return r, best[:n], outc[:n]


def search_by_sim3(fv1, fv2, d1, u12, v12, l12, a1, d2, u21, v21, l21, a2, th, pre12=None):
    arrs = [_u8(d1), _f32(u12), _f32(v12), _i32(l12), _u8(a1), _u8(d2), _f32(u21), _f32(v21), _i32(l21), _u8(a2)]
    m = np.empty(max(fv1.n, 1), np.int32)
    pre = _i32(pre12)
    r = lib().ref_search_by_sim3(C.byref(fv1), C.byref(fv2), *[_p(a) for a in arrs], th, None if pre is None else _p(pre), _p(m))
    return r, m[:fv1.n]


def search_for_triangulation(fv1, has_mp1, featvec1, fv2, has_mp2, featvec2, pose1, only_stereo, coarse, check_ori=True):
    """pose1 = T1w (12 floats: R row-major, t); KF2 at the identity.  Returns (nmatches, match12[n1], ep[2], F12[9])
    where ep and F12 are what the reference's prologue / Pinhole::epipolarConstrain computed from the pose."""
    n1, nd1, o1, i1 = flatten_feature_vector(featvec1)
    n2, nd2, o2, i2 = flatten_feature_vector(featvec2)
    h1 = _u8(has_mp1); h2 = _u8(has_mp2); pose = _f32(pose1)
    m = np.empty(max(fv1.n, 1), np.int32); ep = np.empty(2, np.float32); F = np.empty(9, np.float32)
    r = lib().ref_search_for_triangulation(C.byref(fv1), _p(h1), n1, _p(nd1), _p(o1), _p(i1), C.byref(fv2), _p(h2), n2,
                                           _p(nd2), _p(o2), _p(i2), _p(pose), int(only_stereo), int(coarse),
                                           int(check_ori), _p(ep), _p(F), _p(m))
    return r, m[:fv1.n], ep, F


# ---- two-camera (Frame::Nleft != -1) branches (ref_compat/ref_capi_rig.cpp) ----
def search_by_projection_rig(fvl, fvr, l2r, r2l, mp_desc, px, py, level, view_cos, pxr, pyr, level_r, view_cos_r, depth, flags,
                             th, far_points, th_far, nn_ratio):
    """Returns (nmatches, binding[Nleft + Nright], per_mp[M, 4]): MapPoint bound to each rig feature on return (-1 none,
    -2 pre-existing); the rig features pointing at MapPoint i right after it was processed (ascending, -1 padded)."""
    arrs = [_u8(mp_desc), _f32(px), _f32(py), _i32(level), _f32(view_cos), _f32(pxr), _f32(pyr), _i32(level_r),
            _f32(view_cos_r), _f32(depth), _u8(flags)]
    a, b = _i32(l2r), _i32(r2l)
    out = np.empty(max(fvl.n + fvr.n, 1), np.int32)
    per = np.empty((max(len(arrs[1]), 1), 4), np.int32)
    n = lib().ref_search_by_projection_rig(C.byref(fvl), C.byref(fvr), _p(a), _p(b), len(arrs[1]), *[_p(x) for x in arrs], th,
                                           int(far_points), th_far, nn_ratio, _p(out), _p(per))
    return n, out[:fvl.n + fvr.n], per[:len(arrs[1])]


def search_by_projection_frame_rig(fvl, fvr, last_desc, u, v, z, trl_x, trl_y, last_octave, last_angle, flags, th, forward,
                                   backward, check_ori=True):
    arrs = [_u8(last_desc), _f32(u), _f32(v), _f32(z)]
    tail = [_i32(last_octave), _f32(last_angle), _u8(flags)]
    out = np.empty(max(fvl.n + fvr.n, 1), np.int32)
    n = lib().ref_search_by_projection_frame_rig(C.byref(fvl), C.byref(fvr), len(arrs[1]), *[_p(x) for x in arrs], trl_x, trl_y,
                                                 *[_p(x) for x in tail], th, int(forward), int(backward), int(check_ori), _p(out))
    return n, out[:fvl.n + fvr.n]


def fisheye_stereo(kps_l, desc_l, mono_left, kps_r, desc_r, mono_right, level_sigma2):
    """Frame::ComputeStereoFishEyeMatches (the reference's own body; triangulation = oracle/ref_compat/fake_triangulate.h).
    Returns (mvLeftToRightMatch, mvRightToLeftMatch, mvDepth, mvStereo3Dpoints)."""
    kl = np.ascontiguousarray(kps_l, KP_DTYPE); kr = np.ascontiguousarray(kps_r, KP_DTYPE)
    dl = _u8(desc_l); dr = _u8(desc_r); s2 = _f32(level_sigma2)
    nl, nr = len(kl), len(kr)
    l2r = np.empty(max(nl, 1), np.int32); r2l = np.empty(max(nr, 1), np.int32)
    depth = np.empty(max(nl, 1), np.float32); p3d = np.zeros((max(nl, 1), 3), np.float32)
    lib().ref_fisheye_stereo(nl, _p(kl), _p(dl), int(mono_left), nr, _p(kr), _p(dr), int(mono_right), len(s2), _p(s2), _p(l2r),
                             _p(r2l), _p(depth), _p(p3d))
    return l2r[:nl], r2l[:nr], depth[:nl], p3d[:nl]


def search_by_bow_rig(kl, kr, has_mp1, featvec1, fl, fr, featvec2, nnratio, check_ori=True):
    """ORBmatcher::SearchByBoW(KeyFrame, Frame) with a rig KeyFrame (views kl, kr) and a rig Frame (views fl, fr); the
    FeatureVectors index the combined left-then-right ranges.  Returns (nmatches, match_l[n1], match_r[n1])."""
    n1, nd1, o1, i1 = flatten_feature_vector(featvec1)
    n2, nd2, o2, i2 = flatten_feature_vector(featvec2)
    h1 = _u8(has_mp1)
    N1 = kl.n + kr.n
    ml = np.empty(max(N1, 1), np.int32); mr = np.empty(max(N1, 1), np.int32)
    r = lib().ref_search_by_bow_rig(C.byref(kl), C.byref(kr), _p(h1), n1, _p(nd1), _p(o1), _p(i1), C.byref(fl), C.byref(fr), n2,
                                    _p(nd2), _p(o2), _p(i2), nnratio, int(check_ori), _p(ml), _p(mr))
    return r, ml[:N1], mr[:N1]


def fuse_rig(kl, kr, right, mp_desc, u, v, z, level, active, th):
    """ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight) on a rig KeyFrame (views kl, kr).  Returns (nFused, best_idx[n]) with
    best_idx in the KeyFrame's combined left-then-right index range."""
    arrs = [_u8(mp_desc), _f32(u), _f32(v), _f32(z), _i32(level), _u8(active)]
    n = len(arrs[1])
    best = np.empty(max(n, 1), np.int32)
    r = lib().ref_fuse_rig(C.byref(kl), C.byref(kr), int(right), n, *[_p(a) for a in arrs], th, _p(best))
    return r, best[:n]
