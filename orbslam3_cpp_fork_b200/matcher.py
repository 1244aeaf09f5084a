"""Host-side mirror of the tracking-time subset of class ORB_SLAM3::ORBmatcher (reference
include/ORBmatcher.h:35-158) and of Frame::ComputeStereoMatches (include/Frame.h:155), over the
libcuorb C ABI.  `Frame` / `MapPointSoA` are the flattened views of the reference's pointer graph
that its C++ shim would build.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import KP_DTYPE, check, ptr

FRAME_GRID_ROWS, FRAME_GRID_COLS = 48, 64  # include/Frame.h:36-37


@dataclass
class Frame:
    """The Frame members the matchers read (pinhole / Nleft == -1)."""
    mvKeysUn: np.ndarray            # KP_DTYPE [N]
    mDescriptors: np.ndarray        # u8 [N, 32]
    mvScaleFactors: np.ndarray      # f32 [nlevels]
    width: int
    height: int
    mvuRight: np.ndarray | None = None     # f32 [N], -1 = none
    mb: float = 0.0
    mbf: float = 0.0
    mnMinX: float = 0.0
    mnMinY: float = 0.0
    mnMaxX: float | None = None
    mnMaxY: float | None = None
    # mvpMapPoints: index of the bound MapPoint (-1 = NULL) and whether it has Observations() > 0
    mvpMapPoints: np.ndarray = field(default=None)
    mvpHasObs: np.ndarray = field(default=None)

    def __post_init__(self):
        self.mvKeysUn = np.ascontiguousarray(self.mvKeysUn, KP_DTYPE)
        self.mDescriptors = np.ascontiguousarray(self.mDescriptors, np.uint8)
        self.mvScaleFactors = np.ascontiguousarray(self.mvScaleFactors, np.float32)
        self.N = len(self.mvKeysUn)
        if self.mnMaxX is None: self.mnMaxX = float(self.width)
        if self.mnMaxY is None: self.mnMaxY = float(self.height)
        if self.mvpMapPoints is None: self.mvpMapPoints = np.full(self.N, -1, np.int64)
        if self.mvpHasObs is None: self.mvpHasObs = np.zeros(self.N, np.uint8)
        if self.mvuRight is not None:
            self.mvuRight = np.ascontiguousarray(self.mvuRight, np.float32)

    def view(self, bound=None):
        """cuorb_frame_view + keepalive (src/Frame.cc:240-257 for the grid constants).  bound: the per-feature
        "already taken" flags of the search at hand (default: bound to a MapPoint with observations)."""
        f32 = np.float32
        if bound is None:
            bound = (self.mvpMapPoints >= 0) & (self.mvpHasObs > 0)
        bound = np.ascontiguousarray(np.asarray(bound).astype(np.uint8))
        fv = _lib.FrameView()
        fv.n = self.N
        fv.keys_un = ptr(self.mvKeysUn); fv.descriptors = ptr(self.mDescriptors)
        fv.u_right = ptr(self.mvuRight); fv.bound_with_obs = ptr(bound)
        fv.min_x, fv.min_y, fv.max_x, fv.max_y = self.mnMinX, self.mnMinY, self.mnMaxX, self.mnMaxY
        fv.grid_w_inv = float(f32(FRAME_GRID_COLS) / f32(f32(self.mnMaxX) - f32(self.mnMinX)))
        fv.grid_h_inv = float(f32(FRAME_GRID_ROWS) / f32(f32(self.mnMaxY) - f32(self.mnMinY)))
        fv.mb, fv.mbf = self.mb, self.mbf
        fv.nlevels = len(self.mvScaleFactors)
        fv.scale_factors = ptr(self.mvScaleFactors)
        return fv, bound


@dataclass
class MapPointSoA:
    """Per-MapPoint fields SearchByProjection reads (src/ORBmatcher.cc:50-111), flattened."""
    descriptor: np.ndarray     # u8 [M, 32]   GetDescriptor()
    mTrackProjX: np.ndarray    # f32 [M]
    mTrackProjY: np.ndarray
    mTrackProjXR: np.ndarray
    mnTrackScaleLevel: np.ndarray  # i32 [M]
    mTrackViewCos: np.ndarray
    mTrackDepth: np.ndarray
    mbTrackInView: np.ndarray  # bool [M]
    isBad: np.ndarray          # bool [M]
    hasObservations: np.ndarray  # bool [M]  Observations() > 0

    def flags(self):
        return (np.asarray(self.mbTrackInView, np.uint8) * _lib.MP_IN_VIEW |
                np.asarray(self.isBad, np.uint8) * _lib.MP_BAD |
                np.asarray(self.hasObservations, np.uint8) * _lib.MP_HAS_OBS).astype(np.uint8)


def _c(a, dt):
    return np.ascontiguousarray(a, dt)


class ORBmatcher:
    TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30  # src/ORBmatcher.cc:30-32

    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self._L = _lib.load()
        self.mfNNratio, self.mbCheckOrientation, self.device = float(nnratio), bool(checkOri), int(device)

    # static int DescriptorDistance(const cv::Mat&, const cv::Mat&) — include/ORBmatcher.h:40
    @staticmethod
    def DescriptorDistance(a, b) -> int:
        a = _c(a, np.uint8); b = _c(b, np.uint8)
        assert a.size == 32 and b.size == 32
        return _lib.load().cuorb_hamming256(ptr(a), ptr(b))

    def DescriptorDistanceBatch(self, a, b, ia=None, ib=None):
        a = _c(a, np.uint8).reshape(-1, 32); b = _c(b, np.uint8).reshape(-1, 32)
        ia = None if ia is None else _c(ia, np.int32); ib = None if ib is None else _c(ib, np.int32)
        n = len(ia) if ia is not None else (len(ib) if ib is not None else min(len(a), len(b)))
        out = np.zeros(n, np.int32)
        check(self._L.cuorb_hamming_batch(self.device, ptr(a), len(a), ptr(b), len(b), ptr(ia), ptr(ib), n, ptr(out)))
        return out

    # Frame::ComputeStereoMatches (src/Frame.cc:944-1113); "SearchStereo" in BASELINE.json
    def SearchStereo(self, extractorLeft, extractorRight, mb, mbf, debug=False):
        """Device-resident: matches the last batch of the two extractors.  Returns
        (mvuRight[n, cap], mvDepth[n, cap]) (+ dict of debug arrays)."""
        check(self._L.cuorb_stereo_match_batch_async(extractorLeft.handle, extractorRight.handle, mb, mbf))
        n, cap = extractorLeft._n_last, extractorLeft.capacity
        u = np.zeros((n, cap), np.float32); d = np.zeros((n, cap), np.float32)
        if not debug:
            check(self._L.cuorb_stereo_fetch(extractorLeft.handle, ptr(u), ptr(d), cap, None, None, None))
            return u, d
        bi, bd, sad = (np.zeros((n, cap), np.int32) for _ in range(3))
        check(self._L.cuorb_stereo_fetch(extractorLeft.handle, ptr(u), ptr(d), cap, ptr(bi), ptr(bd), ptr(sad)))
        return u, d, dict(best_idx_r=bi, best_dist=bd, sad=sad)

    def ComputeStereoMatches(self, extractorLeft, extractorRight, mvKeys, mDescriptors, mvKeysRight,
                             mDescriptorsRight, mb, mbf):
        """Host-pointer form with the member set the reference method reads. Returns (mvuRight, mvDepth)."""
        kl = _c(mvKeys, KP_DTYPE); kr = _c(mvKeysRight, KP_DTYPE)
        dl = _c(mDescriptors, np.uint8); dr = _c(mDescriptorsRight, np.uint8)
        u = np.zeros(len(kl), np.float32); d = np.zeros(len(kl), np.float32)
        check(self._L.cuorb_stereo_match(extractorLeft.handle, extractorRight.handle, len(kl), ptr(kl), ptr(dl),
                                         len(kr), ptr(kr), ptr(dr), mb, mbf, ptr(u), ptr(d)))
        return u, d

    def ExtractStereo(self, extractorLeft, extractorRight, imLeft, imRight, mb, mbf):
        """The stereo Frame constructor's three steps as ONE call (cuorb_extract_stereo): ExtractORB on both eyes
        (src/Frame.cc:202-205) + ComputeStereoMatches (:231).  Both extractors must have max_batch == 1.
        Returns (monoIndexLeft, mvKeys, mDescriptors, mvKeysRight, mDescriptorsRight, mvuRight, mvDepth)."""
        imL = np.ascontiguousarray(imLeft, np.uint8); imR = np.ascontiguousarray(imRight, np.uint8)
        if imL.shape != imR.shape or imL.ndim != 2:
            raise ValueError("imLeft / imRight must be two 8-bit single-channel images of one size")
        h, w = imL.shape
        cap = extractorLeft.capacity
        kl = np.zeros(cap, KP_DTYPE); kr = np.zeros(cap, KP_DTYPE)
        dl = np.zeros((cap, 32), np.uint8); dr = np.zeros((cap, 32), np.uint8)
        u = np.zeros(cap, np.float32); d = np.zeros(cap, np.float32)
        nl, nr = C.c_int(0), C.c_int(0)
        mono = check(self._L.cuorb_extract_stereo(extractorLeft.handle, extractorRight.handle, ptr(imL), ptr(imR), w, h, w,
                                                  mb, mbf, ptr(kl), ptr(dl), C.byref(nl), ptr(kr), ptr(dr), C.byref(nr), cap,
                                                  ptr(u), ptr(d)))
        extractorLeft._n_last = extractorRight._n_last = 1
        return (mono, kl[:nl.value].copy(), dl[:nl.value].copy(), kr[:nr.value].copy(), dr[:nr.value].copy(),
                u[:nl.value].copy(), d[:nl.value].copy())

    # int SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints)
    def SearchByProjection(self, F: Frame, mps: MapPointSoA, th=3.0, bFarPoints=False, thFarPoints=50.0):
        fv, keep = F.view()
        M = len(mps.mTrackProjX)
        flags = mps.flags()
        arrs = [_c(mps.descriptor, np.uint8), _c(mps.mTrackProjX, np.float32), _c(mps.mTrackProjY, np.float32),
                _c(mps.mTrackProjXR, np.float32), _c(mps.mnTrackScaleLevel, np.int32),
                _c(mps.mTrackViewCos, np.float32), _c(mps.mTrackDepth, np.float32), flags]
        best = np.zeros(M, np.int32)
        n = check(self._L.cuorb_search_by_projection(self.device, C.byref(fv), M, *[ptr(a) for a in arrs], th,
                                                     int(bFarPoints), thFarPoints, self.mfNNratio, ptr(best)))
        # replay F.mvpMapPoints[bestIdx] = pMP in MapPoint order (src/ORBmatcher.cc:137)
        for i in np.nonzero(best >= 0)[0]:
            F.mvpMapPoints[best[i]] = i
            F.mvpHasObs[best[i]] = 1 if (flags[i] & _lib.MP_HAS_OBS) else 0
        return n, best

    # int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono) with the
    # projections of the last frame's MapPoints supplied by the caller (src/ORBmatcher.cc:1710-1721)
    def SearchByProjectionLastFrame(self, Cur: Frame, last_desc, u, v, inv_z, last_octave, last_angle,
                                    flags, th, bForward=False, bBackward=False):
        fv, keep = Cur.view()
        n_last = len(u)
        flags = _c(flags, np.uint8)
        arrs = [_c(last_desc, np.uint8), _c(u, np.float32), _c(v, np.float32), _c(inv_z, np.float32),
                _c(last_octave, np.int32), _c(last_angle, np.float32), flags]
        best = np.zeros(n_last, np.int32); rot = np.zeros(n_last, np.int32); keepb = np.zeros(3, np.int32)
        n = check(self._L.cuorb_search_by_projection_frame(
            self.device, C.byref(fv), n_last, *[ptr(a) for a in arrs], th, int(bForward), int(bBackward),
            int(self.mbCheckOrientation), ptr(best), ptr(rot), ptr(keepb)))
        for i in np.nonzero(best >= 0)[0]:           # bind in order (:1790)
            Cur.mvpMapPoints[best[i]] = i
            Cur.mvpHasObs[best[i]] = 1 if (flags[i] & _lib.MP_HAS_OBS) else 0
        if self.mbCheckOrientation:                   # rotation consistency (:1889-1905)
            for i in np.nonzero(best >= 0)[0]:
                if rot[i] not in keepb:
                    Cur.mvpMapPoints[best[i]] = -1
        return n, best, rot, keepb


@dataclass
class RigFrame:
    """A two-camera Frame (Nleft != -1): the left camera's features are indices [0, Nleft) of the reference's arrays
    (mvKeys, mGrid), the right camera's [Nleft, N) (mvKeysRight, mGridRight, descriptor rows idx + Nleft).  Held as two
    Frame mirrors that share bounds / scale factors; mvpMapPoints of the reference = left.mvpMapPoints followed by
    right.mvpMapPoints."""
    left: Frame
    right: Frame
    mvLeftToRightMatch: np.ndarray      # i32 [Nleft],  -1 = no stereo partner
    mvRightToLeftMatch: np.ndarray      # i32 [Nright]

    def __post_init__(self):
        self.mvLeftToRightMatch = np.ascontiguousarray(self.mvLeftToRightMatch, np.int32)
        self.mvRightToLeftMatch = np.ascontiguousarray(self.mvRightToLeftMatch, np.int32)
        self.Nleft, self.Nright = self.left.N, self.right.N

    def _write(self, side, idx, who, has_obs):
        f = self.right if side else self.left
        f.mvpMapPoints[idx] = who
        f.mvpHasObs[idx] = 1 if has_obs else 0


def SearchByProjectionRig(self, F: RigFrame, mp_desc, px, py, level, view_cos, pxr, pyr, level_r, view_cos_r, depth, flags,
                          th=3.0, bFarPoints=False, thFarPoints=50.0):
    """int SearchByProjection(Frame&, vpMapPoints, th, bFarPoints, thFarPoints) with F.Nleft != -1
    (src/ORBmatcher.cc:38-227).  Returns (nmatches, best_l, best_r) and replays the reference's writes into
    F.left / F.right mvpMapPoints in MapPoint order, stereo partners included (:138-142, :212-216)."""
    fvl, keep_l = F.left.view(); fvr, keep_r = F.right.view()
    flags = _c(flags, np.uint8)
    arrs = [_c(mp_desc, np.uint8), _c(px, np.float32), _c(py, np.float32), _c(level, np.int32), _c(view_cos, np.float32),
            _c(pxr, np.float32), _c(pyr, np.float32), _c(level_r, np.int32), _c(view_cos_r, np.float32), _c(depth, np.float32), flags]
    M = len(arrs[1])
    bl = np.full(max(M, 1), -1, np.int32); br = np.full(max(M, 1), -1, np.int32)
    n = check(self._L.cuorb_search_by_projection_rig(self.device, C.byref(fvl), C.byref(fvr), ptr(F.mvLeftToRightMatch),
                                                     ptr(F.mvRightToLeftMatch), M, *[ptr(a) for a in arrs], th, int(bFarPoints),
                                                     thFarPoints, self.mfNNratio, ptr(bl), ptr(br)))
    bl, br = bl[:M], br[:M]
    for i in np.nonzero((bl >= 0) | (br >= 0))[0]:
        obs = bool(flags[i] & _lib.MP_HAS_OBS)
        if bl[i] >= 0:
            F._write(0, bl[i], i, obs)
            if F.mvLeftToRightMatch[bl[i]] != -1:
                F._write(1, F.mvLeftToRightMatch[bl[i]], i, obs)
        if br[i] >= 0:
            if F.mvRightToLeftMatch[br[i]] != -1:
                F._write(0, F.mvRightToLeftMatch[br[i]], i, obs)
            F._write(1, br[i], i, obs)
    return n, bl, br


def SearchByProjectionLastFrameRig(self, Cur: RigFrame, last_desc, u, v, inv_z, ur, vr, last_octave, last_angle, flags, th,
                                   bForward=False, bBackward=False):
    """int SearchByProjection(Frame& Cur, const Frame& Last, th, bMono) with CurrentFrame.Nleft != -1
    (src/ORBmatcher.cc:1684-1908); (ur, vr) = project(Trl * x3Dc).  Returns (nmatches, best_l, best_r, rot_l, rot_r, keep)."""
    fvl, keep_l = Cur.left.view(); fvr, keep_r = Cur.right.view()
    flags = _c(flags, np.uint8)
    arrs = [_c(last_desc, np.uint8), _c(u, np.float32), _c(v, np.float32), _c(inv_z, np.float32), _c(ur, np.float32),
            _c(vr, np.float32), _c(last_octave, np.int32), _c(last_angle, np.float32), flags]
    n_last = len(arrs[1])
    outs = [np.full(max(n_last, 1), -1, np.int32) for _ in range(4)] + [np.full(3, -1, np.int32)]
    n = check(self._L.cuorb_search_by_projection_frame_rig(self.device, C.byref(fvl), C.byref(fvr), n_last, *[ptr(a) for a in arrs],
                                                           th, int(bForward), int(bBackward), int(self.mbCheckOrientation),
                                                           *[ptr(o) for o in outs]))
    bl, br, rl, rr = (o[:n_last] for o in outs[:4])
    keepb = outs[4]
    for i in range(n_last):
        obs = bool(flags[i] & _lib.MP_HAS_OBS)
        if bl[i] >= 0: Cur._write(0, bl[i], i, obs)
        if br[i] >= 0: Cur._write(1, br[i], i, obs)
    if self.mbCheckOrientation:
        for i in range(n_last):
            if bl[i] >= 0 and rl[i] not in keepb: Cur.left.mvpMapPoints[bl[i]] = -1
            if br[i] >= 0 and rr[i] not in keepb: Cur.right.mvpMapPoints[br[i]] = -1
    return n, bl, br, rl, rr, keepb


ORBmatcher.SearchByProjectionRig = SearchByProjectionRig
ORBmatcher.SearchByProjectionLastFrameRig = SearchByProjectionLastFrameRig


def join_feature_vectors(mFeatVec1, mFeatVec2):
    """The merge-join SearchByBoW does over two DBoW2::FeatureVectors (std::map<NodeId, std::vector<unsigned>>,
    src/ORBmatcher.cc:252-259, 395-399): common nodes in ascending NodeId order.
    mFeatVec*: dict {node_id: sequence of feature indices}.  Returns (off1, idx1, off2, idx2)."""
    common = sorted(set(mFeatVec1) & set(mFeatVec2))
    off1 = np.zeros(len(common) + 1, np.int32); off2 = np.zeros(len(common) + 1, np.int32)
    l1, l2 = [], []
    for k, node in enumerate(common):
        a = np.asarray(mFeatVec1[node], np.int32); b = np.asarray(mFeatVec2[node], np.int32)
        l1.append(a); l2.append(b)
        off1[k + 1] = off1[k] + len(a); off2[k + 1] = off2[k] + len(b)
    idx1 = np.concatenate(l1).astype(np.int32) if l1 else np.zeros(0, np.int32)
    idx2 = np.concatenate(l2).astype(np.int32) if l2 else np.zeros(0, np.int32)
    return off1, idx1, off2, idx2


def _search_by_bow(self, desc1, valid1, angle1, mFeatVec1, desc2, eligible2, angle2, mFeatVec2, mode):
    desc1 = _c(desc1, np.uint8).reshape(-1, 32); desc2 = _c(desc2, np.uint8).reshape(-1, 32)
    valid1 = _c(np.asarray(valid1).astype(np.uint8), np.uint8); eligible2 = _c(np.asarray(eligible2).astype(np.uint8), np.uint8)
    angle1 = _c(angle1, np.float32); angle2 = _c(angle2, np.float32)
    off1, idx1, off2, idx2 = join_feature_vectors(mFeatVec1, mFeatVec2)
    match12 = np.full(len(desc1), -1, np.int32)
    n = check(self._L.cuorb_search_by_bow(self.device, len(desc1), ptr(desc1), ptr(valid1), ptr(angle1), len(desc2), ptr(desc2),
                                          ptr(eligible2), ptr(angle2), len(off1) - 1, ptr(off1), ptr(idx1), ptr(off2), ptr(idx2),
                                          int(mode), self.mfNNratio, int(self.mbCheckOrientation), ptr(match12)))
    return n, match12


def SearchByBoW_KF_F(self, kf_desc, kf_has_mappoint, kf_angle, kf_featvec, f_desc, f_angle, f_featvec):
    """int SearchByBoW(KeyFrame* pKF, Frame& F, vpMapPointMatches) (src/ORBmatcher.cc:237-420), mono branch.
    Returns (nmatches, match[kf feature] = F feature or -1); the reference's
    vpMapPointMatches[match[i]] = pKF->GetMapPointMatches()[i] is the caller's replay."""
    return _search_by_bow(self, kf_desc, kf_has_mappoint, kf_angle, kf_featvec, f_desc, np.ones(len(f_desc), np.uint8), f_angle,
                          f_featvec, 0)


def SearchByBoW_KF_KF(self, desc1, has_mp1, angle1, featvec1, desc2, has_mp2, angle2, featvec2):
    """int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12) (src/ORBmatcher.cc:784-916).
    Returns (nmatches, match12[i1] = feature of pKF2 or -1)."""
    return _search_by_bow(self, desc1, has_mp1, angle1, featvec1, desc2, has_mp2, angle2, featvec2, 1)


def SearchByBoW_KF_RigF(self, kf_desc, kf_has_mappoint, kf_angle, kf_featvec, f_desc, f_nleft, f_angle, f_featvec):
    """int SearchByBoW(KeyFrame* pKF, Frame& F, vpMapPointMatches) with F.Nleft != -1 (src/ORBmatcher.cc:306-331, :371-396).
    f_desc / f_angle cover F's left-then-right features, f_nleft = F.Nleft.  Returns (nmatches, match_left, match_right)."""
    desc1 = _c(kf_desc, np.uint8).reshape(-1, 32); desc2 = _c(f_desc, np.uint8).reshape(-1, 32)
    valid1 = _c(np.asarray(kf_has_mappoint).astype(np.uint8), np.uint8)
    angle1 = _c(kf_angle, np.float32); angle2 = _c(f_angle, np.float32)
    off1, idx1, off2, idx2 = join_feature_vectors(kf_featvec, f_featvec)
    ml = np.full(max(len(desc1), 1), -1, np.int32); mr = np.full(max(len(desc1), 1), -1, np.int32)
    n = check(self._L.cuorb_search_by_bow_rig(self.device, len(desc1), ptr(desc1), ptr(valid1), ptr(angle1), len(desc2), int(f_nleft),
                                              ptr(desc2), ptr(angle2), len(off1) - 1, ptr(off1), ptr(idx1), ptr(off2), ptr(idx2),
                                              self.mfNNratio, int(self.mbCheckOrientation), ptr(ml), ptr(mr)))
    return n, ml[:len(desc1)], mr[:len(desc1)]


ORBmatcher.SearchByBoW_KF_F = SearchByBoW_KF_F
ORBmatcher.SearchByBoW_KF_RigF = SearchByBoW_KF_RigF
ORBmatcher.SearchByBoW_KF_KF = SearchByBoW_KF_KF


# ---- the remaining searches; a KeyFrame is passed as a Frame (same members) ----
def SearchByProjectionReloc(self, Cur: Frame, mp_desc, u, v, level, kf_angle, active, th, ORBdist):
    """int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, sAlreadyFound, th, ORBdist)
    (src/ORBmatcher.cc:1910-2036).  Binds Cur.mvpMapPoints like the reference; returns (nmatches, best, rot, keep)."""
    fv, keep = Cur.view(bound=Cur.mvpMapPoints >= 0)
    n = len(u)
    arrs = [_c(mp_desc, np.uint8), _c(u, np.float32), _c(v, np.float32), _c(level, np.int32), _c(kf_angle, np.float32),
            _c(np.asarray(active).astype(np.uint8), np.uint8)]
    best = np.zeros(n, np.int32); rot = np.zeros(n, np.int32); keepb = np.zeros(3, np.int32)
    r = check(self._L.cuorb_search_by_projection_reloc(self.device, C.byref(fv), n, *[ptr(a) for a in arrs], th, int(ORBdist),
                                                       int(self.mbCheckOrientation), ptr(best), ptr(rot), ptr(keepb)))
    for i in np.nonzero(best >= 0)[0]:
        Cur.mvpMapPoints[best[i]] = i if (not self.mbCheckOrientation or rot[i] in keepb) else -1
    return r, best, rot, keepb


def SearchByProjectionSim3(self, KF: Frame, vpMatched_set, mp_desc, u, v, level, active, th, ratioHamming):
    """int SearchByProjection(KeyFrame* pKF, Sim3& Scw, vpPoints, [vpPointsKFs,] vpMatched, [vpMatchedKF,] th,
    ratioHamming) (src/ORBmatcher.cc:423-538, 540-664).  vpMatched_set[idx] = vpMatched[idx] != NULL on entry."""
    fv, keep = KF.view(bound=vpMatched_set)
    n = len(u)
    arrs = [_c(mp_desc, np.uint8), _c(u, np.float32), _c(v, np.float32), _c(level, np.int32),
            _c(np.asarray(active).astype(np.uint8), np.uint8)]
    best = np.zeros(n, np.int32)
    r = check(self._L.cuorb_search_by_projection_sim3(self.device, C.byref(fv), n, *[ptr(a) for a in arrs], int(th),
                                                      ratioHamming, ptr(best)))
    return r, best


def Fuse(self, KF: Frame, mvInvLevelSigma2, mp_desc, u, v, ur, level, active, th=3.0, check_reprojection=True):
    """int Fuse(KeyFrame*, vpMapPoints, th, bRight=false) (src/ORBmatcher.cc:1149-1328; check_reprojection=True) and
    int Fuse(KeyFrame*, Sim3f&, vpPoints, th, vpReplacePoint) (:1330-1452; check_reprojection=False)."""
    fv, keep = KF.view()
    n = len(u)
    inv = _c(mvInvLevelSigma2, np.float32) if mvInvLevelSigma2 is not None else None
    arrs = [_c(mp_desc, np.uint8), _c(u, np.float32), _c(v, np.float32), None if ur is None else _c(ur, np.float32),
            _c(level, np.int32), _c(np.asarray(active).astype(np.uint8), np.uint8)]
    best = np.zeros(n, np.int32)
    r = check(self._L.cuorb_fuse(self.device, C.byref(fv), ptr(inv), n, *[ptr(a) for a in arrs], th,
                                 int(check_reprojection), ptr(best)))
    return r, best


def FuseRig(self, KF: "RigFrame", bRight, mvInvLevelSigma2, mp_desc, u, v, ur, level, active, th=3.0, check_reprojection=True):
    """Fuse on a two-camera rig KeyFrame (NLeft != -1; src/ORBmatcher.cc:1156-1164, :1236, :1270-1272): the search runs on
    the left camera's features or, with bRight, on the right camera's; best indices come back in the KeyFrame's combined
    left-then-right range, as the reference leaves them."""
    side = KF.right if bRight else KF.left
    saved = side.mvuRight
    side.mvuRight = None                                           # a rig KeyFrame's mvuRight is all -1
    try:
        r, best = Fuse(self, side, mvInvLevelSigma2, mp_desc, u, v, ur, level, active, th, check_reprojection)
    finally:
        side.mvuRight = saved
    if bRight:
        best = np.where(best >= 0, best + KF.Nleft, -1).astype(np.int32)
    return r, best


def SearchBySim3(self, KF1: Frame, KF2: Frame, mp_desc1, u12, v12, level12, active1, mp_desc2, u21, v21, level21,
                 active2, th):
    """int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12, S12, th) (src/ORBmatcher.cc:1454-1682)."""
    fv1, k1 = KF1.view(); fv2, k2 = KF2.view()
    arrs = [_c(mp_desc1, np.uint8), _c(u12, np.float32), _c(v12, np.float32), _c(level12, np.int32),
            _c(np.asarray(active1).astype(np.uint8), np.uint8), _c(mp_desc2, np.uint8), _c(u21, np.float32),
            _c(v21, np.float32), _c(level21, np.int32), _c(np.asarray(active2).astype(np.uint8), np.uint8)]
    m12 = np.full(max(KF1.N, 1), -1, np.int32)
    r = check(self._L.cuorb_search_by_sim3(self.device, C.byref(fv1), C.byref(fv2), *[ptr(a) for a in arrs], th, ptr(m12)))
    return r, m12[:KF1.N]


def SearchForInitialization(self, F1: Frame, F2: Frame, vbPrevMatched, windowSize=10):
    """int SearchForInitialization(Frame& F1, Frame& F2, vbPrevMatched, vnMatches12, windowSize)
    (src/ORBmatcher.cc:666-782).  vbPrevMatched: f32 [N1, 2], updated in place.  Returns (nmatches, vnMatches12)."""
    fv1, k1 = F1.view(); fv2, k2 = F2.view()
    pm = _c(vbPrevMatched, np.float32).reshape(-1, 2).copy()
    m12 = np.full(max(F1.N, 1), -1, np.int32)
    r = check(self._L.cuorb_search_for_initialization(self.device, C.byref(fv1), C.byref(fv2), ptr(pm), int(windowSize),
                                                      self.mfNNratio, int(self.mbCheckOrientation), ptr(m12)))
    np.asarray(vbPrevMatched)[...] = pm.reshape(np.asarray(vbPrevMatched).shape)
    return r, m12[:F1.N]


def SearchForTriangulation(self, KF1: Frame, KF2: Frame, has_mp1, has_mp2, featvec1, featvec2, F12, epipole,
                           mvLevelSigma2_2, bOnlyStereo=False, bCoarse=False):
    """int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, vMatchedPairs, bOnlyStereo, bCoarse)
    (src/ORBmatcher.cc:916-1147).  Returns (nmatches, match12); vMatchedPairs = [(i, match12[i]) for match12[i] >= 0]."""
    fv1, k1 = KF1.view(bound=has_mp1); fv2, k2 = KF2.view(bound=has_mp2)
    off1, idx1, off2, idx2 = join_feature_vectors(featvec1, featvec2)
    F12 = _c(np.asarray(F12, np.float32).reshape(9), np.float32); ep = _c(epipole, np.float32)
    s2 = _c(mvLevelSigma2_2, np.float32)
    m12 = np.full(max(KF1.N, 1), -1, np.int32)
    r = check(self._L.cuorb_search_for_triangulation(self.device, C.byref(fv1), C.byref(fv2), len(off1) - 1, ptr(off1),
                                                     ptr(idx1), ptr(off2), ptr(idx2), ptr(F12), ptr(ep), ptr(s2),
                                                     int(bOnlyStereo), int(bCoarse), int(self.mbCheckOrientation), ptr(m12)))
    return r, m12[:KF1.N]


ORBmatcher.SearchByProjectionReloc = SearchByProjectionReloc
ORBmatcher.SearchByProjectionSim3 = SearchByProjectionSim3
ORBmatcher.Fuse = Fuse
ORBmatcher.FuseRig = FuseRig
ORBmatcher.SearchBySim3 = SearchBySim3
ORBmatcher.SearchForInitialization = SearchForInitialization
ORBmatcher.SearchForTriangulation = SearchForTriangulation
