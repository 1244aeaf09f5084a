"""Host-side mirror of the dense all-pairs Hamming users of the reference:
cv::BFMatcher(NORM_HAMMING).knnMatch(..., 2) + Lowe ratio of Frame::ComputeStereoFishEyeMatches
(src/Frame.cc:1304-1355) and MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:376-450).
Compute is libcuorb's (no CPU fallback)."""
from __future__ import annotations

import numpy as np

from . import _lib
from ._lib import check, ptr


def _csr(sets):
    sets = [np.ascontiguousarray(s, dtype=np.uint8).reshape(-1, 32) for s in sets]
    off = np.zeros(len(sets) + 1, np.int32)
    off[1:] = np.cumsum([len(s) for s in sets])
    flat = np.concatenate(sets) if sets else np.zeros((0, 32), np.uint8)
    return off, np.ascontiguousarray(flat)


def knnMatch2(query_sets, train_sets, ratio_test=False, device=0):
    """query_sets / train_sets: lists of [n, 32] uint8 descriptor matrices (one pair per frame).
    Returns per set (idx [nq, 2], dist [nq, 2][, match [nq]]) like BFMatcher.knnMatch(k=2):
    trainIdx / distance of the two nearest rows in (distance, trainIdx) order, -1 where missing."""
    L = _lib.load()
    if len(query_sets) != len(train_sets):
        raise ValueError("need one train set per query set")
    qo, q = _csr(query_sets)
    to, t = _csr(train_sets)
    nq = int(qo[-1])
    idx = np.full((nq, 2), -1, np.int32); dist = np.full((nq, 2), -1, np.int32)
    match = np.full(nq, -1, np.int32) if ratio_test else None
    if nq:
        tt = t if len(t) else np.zeros((1, 32), np.uint8)
        check(L.cuorb_knn_match2(int(device), len(query_sets), ptr(qo), ptr(q), ptr(to), ptr(tt), ptr(idx), ptr(dist), ptr(match)))
    out = []
    for s in range(len(query_sets)):
        a, b = qo[s], qo[s + 1]
        out.append((idx[a:b], dist[a:b]) + ((match[a:b],) if ratio_test else ()))
    return out


def ComputeDistinctiveDescriptors(observed, device=0):
    """observed: list (one entry per MapPoint) of [n_obs, 32] uint8 matrices in the reference's vDescriptors
    order.  Returns BestIdx per MapPoint (-1 for a point without observations)."""
    L = _lib.load()
    off, flat = _csr(observed)
    best = np.full(len(observed), -1, np.int32)
    if len(observed):
        ff = flat if len(flat) else np.zeros((1, 32), np.uint8)
        check(L.cuorb_distinctive_descriptors(int(device), len(observed), ptr(off), ptr(ff), ptr(best)))
    return best
