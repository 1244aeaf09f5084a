"""CPU: oracle restatement of BFMatcher.knnMatch(k=2) + ratio test and of
MapPoint::ComputeDistinctiveDescriptors, against cv2 goldens / a numpy re-derivation."""
import os

import numpy as np
import pytest

from oracle import oracle_py as op

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "bfmatch_cv2.npz")
CASES = ("random", "ties6", "ties3", "two", "one")


@pytest.mark.parametrize("case", CASES)
def test_oracle_knn2_matches_cv2_golden(case):
    g = np.load(GOLDEN)
    idx, dist = op.knn_match2(g[f"{case}_q"], g[f"{case}_t"])
    assert (idx == g[f"{case}_idx"]).all() and (dist == g[f"{case}_dist"]).all()
    assert (op.fisheye_ratio_test(idx, dist) == g[f"{case}_ratio"]).all()


def observations(seed, n_points=40, max_obs=70):
    """Per MapPoint: noisy copies of a base descriptor (tie-prone medians), incl. 0-, 1- and 2-observation points."""
    rng = np.random.default_rng(seed)
    out = []
    for p in range(n_points):
        n = [0, 1, 2, 3][p] if p < 4 else int(rng.integers(2, max_obs))
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        d = np.repeat(base[None], n, 0)
        flips = rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8) & rng.integers(0, 256, (n, 32), dtype=np.uint8)
        out.append(d ^ flips if n else d.reshape(0, 32))
    return out


def distinctive_numpy(d):
    n = len(d)
    if n == 0:
        return -1
    D = np.unpackbits(d[:, None, :] ^ d[None, :, :], axis=2).sum(2)
    med = np.sort(D, axis=1)[:, int(0.5 * (n - 1))]
    return int(np.argmin(med))                                   # first minimum (src/MapPoint.cc:434)


def test_oracle_distinctive_descriptor():
    obs = observations(1)
    off = np.zeros(len(obs) + 1, np.int32); off[1:] = np.cumsum([len(o) for o in obs])
    best = op.distinctive_descriptors(np.concatenate(obs), off)
    assert best.tolist() == [distinctive_numpy(o) for o in obs]
