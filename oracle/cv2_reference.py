"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

The reference's ORBextractor::operator() (src/ORBextractor.cc:1052-1138) re-assembled in Python
around the REAL OpenCV primitives (cv2 4.13.0: resize, copyMakeBorder, FastFeatureDetector,
GaussianBlur, fastAtan2).  It exists to pin the C++ oracle's restated primitives to the real
library and to generate the golden vectors in tests/golden/ (tests/golden/make_golden.py).

The one stage taken from the C++ oracle is DistributeOctTree: its output order depends on
libstdc++'s std::list / non-stable std::sort behaviour, which Python cannot reproduce, so the
oracle's std::list + std::sort implementation (oracle/orb_oracle.cpp) is called for it.
glibc cosf/sinf are called through ctypes so the descriptor rotation uses the same libm as C++.
"""
from __future__ import annotations

import ctypes
import ctypes.util
import math

import cv2
import numpy as np

from . import oracle_py as op

_libm = ctypes.CDLL(ctypes.util.find_library("m"))
_libm.cosf.restype = ctypes.c_float
_libm.cosf.argtypes = [ctypes.c_float]
_libm.sinf.restype = ctypes.c_float
_libm.sinf.argtypes = [ctypes.c_float]

EDGE = 19
F32 = np.float32


def cv_round(v) -> int:
    return int(np.rint(v))


def tables(nfeatures, scale_factor, nlevels):
    """ctor arithmetic (src/ORBextractor.cc:411-440) in float32/double as the C++ does."""
    sf = np.zeros(nlevels, F32)
    sf[0] = 1.0
    sfd = float(F32(scale_factor))  # float ctor arg stored into the double member
    for i in range(1, nlevels):
        sf[i] = F32(float(sf[i - 1]) * sfd)
    inv = (F32(1.0) / sf).astype(F32)
    factor = F32(1.0 / sfd)
    nd = F32(F32(nfeatures) * F32(F32(1) - factor)) / F32(F32(1) - F32(math.pow(float(factor), float(nlevels))))
    nd = F32(nd)
    per = np.zeros(nlevels, np.int32)
    s = 0
    for lv in range(nlevels - 1):
        per[lv] = cv_round(nd)
        s += per[lv]
        nd = F32(nd * factor)
    per[nlevels - 1] = max(nfeatures - s, 0)
    return sf, inv, per


def umax_table():
    hp = 15
    umax = [0] * (hp + 1)
    vmax = int(math.floor(F32(hp) * F32(math.sqrt(2.0)) / 2 + 1))
    vmin = int(math.ceil(F32(hp) * F32(math.sqrt(2.0)) / 2))
    for v in range(vmax + 1):
        umax[v] = cv_round(math.sqrt(hp * hp - v * v))
    v0 = 0
    for v in range(hp, vmin - 1, -1):
        while umax[v0] == umax[v0 + 1]:
            v0 += 1
        umax[v] = v0
        v0 += 1
    return umax


def compute_pyramid(image, inv_sf):
    """ComputePyramid (src/ORBextractor.cc:1140-1173). Returns (wholes, rois) lists."""
    wholes, rois = [], []
    for level, scale in enumerate(inv_sf):
        sw = cv_round(F32(image.shape[1]) * scale)
        sh = cv_round(F32(image.shape[0]) * scale)
        if level != 0:
            lvl = cv2.resize(np.ascontiguousarray(rois[level - 1]), (sw, sh), fx=0, fy=0,
                             interpolation=cv2.INTER_LINEAR)
            temp = cv2.copyMakeBorder(lvl, EDGE, EDGE, EDGE, EDGE,
                                      cv2.BORDER_REFLECT_101 + cv2.BORDER_ISOLATED)
        else:
            temp = cv2.copyMakeBorder(np.ascontiguousarray(image), EDGE, EDGE, EDGE, EDGE,
                                      cv2.BORDER_REFLECT_101)
        assert temp.shape == (sh + 2 * EDGE, sw + 2 * EDGE)
        roi = temp[EDGE:EDGE + sh, EDGE:EDGE + sw]
        wholes.append(temp)
        rois.append(roi)
    return wholes, rois


def fast_cells(level_img, ini_th, min_th):
    """The cell loop of ComputeKeyPointsOctTree (src/ORBextractor.cc:755-850) with cv2 FAST.
    Returns list of (x, y, response) in vToDistributeKeys order and the region bounds."""
    rows, cols = level_img.shape
    min_bx = min_by = EDGE - 3
    max_bx = cols - EDGE + 3
    max_by = rows - EDGE + 3
    width = F32(max_bx - min_bx); height = F32(max_by - min_by)
    W = F32(35)
    n_cols = int(width / W); n_rows = int(height / W)
    w_cell = int(math.ceil(width / F32(n_cols))); h_cell = int(math.ceil(height / F32(n_rows)))
    det_ini = cv2.FastFeatureDetector_create(ini_th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    det_min = cv2.FastFeatureDetector_create(min_th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    out = []
    for i in range(n_rows):
        ini_y = min_by + i * h_cell
        max_y = ini_y + h_cell + 6
        if ini_y >= max_by - 3:
            continue
        max_y = min(max_y, max_by)
        for j in range(n_cols):
            ini_x = min_bx + j * w_cell
            max_x = ini_x + w_cell + 6
            if ini_x >= max_bx - 6:
                continue
            max_x = min(max_x, max_bx)
            roi = level_img[ini_y:max_y, ini_x:max_x]
            kps = det_ini.detect(roi)
            if len(kps) == 0:
                kps = det_min.detect(roi)
            for kp in kps:
                out.append((kp.pt[0] + j * w_cell, kp.pt[1] + i * h_cell, kp.response))
    return out, (min_bx, max_bx, min_by, max_by)


def ic_angle(level_img, x, y, umax):
    """IC_Angle (src/ORBextractor.cc:66-91) with cv2.fastAtan2."""
    m01 = 0
    m10 = 0
    img = level_img
    row = img[y].astype(np.int64)
    us = np.arange(-15, 16)
    m10 += int((us * row[x - 15:x + 16]).sum())
    for v in range(1, 16):
        d = umax[v]
        u = np.arange(-d, d + 1)
        plus = img[y + v, x - d:x + d + 1].astype(np.int64)
        minus = img[y - v, x - d:x + d + 1].astype(np.int64)
        m01 += v * int((plus - minus).sum())
        m10 += int((u * (plus + minus)).sum())
    return cv2.fastAtan2(float(F32(m01)), float(F32(m10)))


_PATTERN = None


def pattern():
    global _PATTERN
    if _PATTERN is None:
        import os, re
        hdr = open(os.path.join(os.path.dirname(__file__), "..", "include", "cuorb_pattern.h")).read()
        body = hdr[hdr.index("{") + 1: hdr.rindex("}")]
        vals = [int(t) for t in re.findall(r"-?\d+", body)]
        assert len(vals) == 1024
        _PATTERN = np.array(vals, np.int32).reshape(256, 2, 2)  # [pair][tap][x,y]
    return _PATTERN


def orb_descriptor(blurred, x, y, angle_deg):
    """computeOrbDescriptor (src/ORBextractor.cc:93-139): float32 math, glibc cosf/sinf,
    cvRound (half-even), no FMA (numpy evaluates the two products and the sum separately)."""
    factor_pi = F32(math.pi / float(F32(180.0)))
    ang = F32(F32(angle_deg) * factor_pi)
    a = F32(_libm.cosf(float(ang)))
    b = F32(_libm.sinf(float(ang)))
    pat = pattern()
    px = pat[:, :, 0].astype(F32); py = pat[:, :, 1].astype(F32)
    dy = np.rint((px * b).astype(F32) + (py * a).astype(F32)).astype(np.int32)
    dx = np.rint((px * a).astype(F32) - (py * b).astype(F32)).astype(np.int32)
    vals = blurred[y + dy, x + dx].astype(np.int32)  # [256, 2]
    bits = (vals[:, 0] < vals[:, 1]).astype(np.uint8).reshape(32, 8)
    return (bits << np.arange(8, dtype=np.uint8)).sum(axis=1).astype(np.uint8)


def extract(image, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, lap=(0, 0),
            return_stages=False):
    """operator() with real cv2 primitives. Returns (monoIndex, keypoints[KP_DTYPE], desc[n,32])."""
    if image is None or image.size == 0:
        return -1, np.zeros(0, op.KP_DTYPE), np.zeros((0, 32), np.uint8)
    sf, inv, per = tables(nfeatures, scale_factor, nlevels)
    umax = umax_table()
    wholes, rois = compute_pyramid(image, inv)
    oct_helper = op.OracleExtractor(nfeatures, scale_factor, nlevels, ini_th, min_th)
    all_keys = []
    stages = dict(pyramid=wholes, candidates=[], level_keys=[], blurred=[])
    for level in range(nlevels):
        cands, (min_bx, max_bx, min_by, max_by) = fast_cells(rois[level], ini_th, min_th)
        stages["candidates"].append(cands)
        keys = np.zeros(len(cands), op.KP_DTYPE)
        for i, (x, y, r) in enumerate(cands):
            keys[i] = (x, y, 7.0, -1.0, r, 0, -1)
        sel = oct_helper.octree(keys, min_bx, max_bx, min_by, max_by, int(per[level])) if len(keys) else keys
        sel["x"] += min_bx
        sel["y"] += min_by
        sel["octave"] = level
        sel["size"] = F32(int(F32(31) * sf[level]))
        all_keys.append(sel)
    for level in range(nlevels):
        for kp in all_keys[level]:
            kp["angle"] = ic_angle(wholes[level], int(kp["x"]) + EDGE, int(kp["y"]) + EDGE, umax)
        stages["level_keys"].append(all_keys[level].copy())
    n = sum(len(k) for k in all_keys)
    out_k = np.zeros(n, op.KP_DTYPE)
    out_d = np.zeros((n, 32), np.uint8)
    mono, stereo = 0, n - 1
    for level in range(nlevels):
        keys = all_keys[level]
        if len(keys) == 0:
            stages["blurred"].append(None)
            continue
        working = rois[level].copy()
        working = cv2.GaussianBlur(working, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        stages["blurred"].append(working)
        scale = sf[level]
        for kp in keys:
            d = orb_descriptor(working, cv_round(kp["x"]), cv_round(kp["y"]), kp["angle"])
            if level != 0:
                kp["x"] = F32(kp["x"] * scale)
                kp["y"] = F32(kp["y"] * scale)
            if lap[0] <= kp["x"] <= lap[1]:
                out_k[stereo] = kp; out_d[stereo] = d; stereo -= 1
            else:
                out_k[mono] = kp; out_d[mono] = d; mono += 1
    if return_stages:
        return mono, out_k, out_d, stages
    return mono, out_k, out_d
