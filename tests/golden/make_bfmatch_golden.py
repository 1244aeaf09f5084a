"""Golden vectors for the brute-force 2-NN Hamming matcher made with the REAL
cv2.BFMatcher(NORM_HAMMING).knnMatch(k=2) (4.13.0), incl. tie-heavy and tiny train sets.
Run here (container with cv2):  python tests/golden/make_bfmatch_golden.py"""
import os

import numpy as np

CASES = {"random": (300, 400, 256), "ties6": (200, 150, 6), "ties3": (257, 513, 3), "two": (50, 2, 256), "one": (64, 1, 256)}


def make_sets(rng, nq, nt, nbits):
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8); t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    if nbits < 256:                       # only a few bits vary: many equal distances
        mask = np.zeros(32, np.uint8); mask[0] = (1 << nbits) - 1
        q &= mask; t &= mask
    return q, t


def main():
    import cv2
    rng = np.random.default_rng(20261106)
    rec = dict(cv2_version=cv2.__version__)
    for name, (nq, nt, nbits) in CASES.items():
        q, t = make_sets(rng, nq, nt, nbits)
        m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, k=2)
        idx = np.full((nq, 2), -1, np.int32); dist = np.full((nq, 2), -1, np.int32)
        for i, mm in enumerate(m):
            for k, x in enumerate(mm):
                idx[i, k] = x.trainIdx; dist[i, k] = int(x.distance)
        ratio = np.array([mm[0].trainIdx if len(mm) >= 2 and mm[0].distance < mm[1].distance * 0.7 else -1 for mm in m], np.int32)
        rec[f"{name}_q"] = q; rec[f"{name}_t"] = t; rec[f"{name}_idx"] = idx; rec[f"{name}_dist"] = dist; rec[f"{name}_ratio"] = ratio
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "bfmatch_cv2.npz"), **rec)
    print("wrote bfmatch_cv2.npz")


if __name__ == "__main__":
    main()
