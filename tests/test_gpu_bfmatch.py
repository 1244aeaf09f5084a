"""GPU: libcuorb's brute-force 2-NN Hamming matcher and distinctive-descriptor selection (through the C
ABI) — bit-exact against the cv2 goldens and the CPU oracle."""
import numpy as np
import pytest

from tests.test_oracle_bfmatch import CASES, GOLDEN, observations

pytestmark = pytest.mark.gpu


def test_knn2_matches_cv2_goldens_batched():
    from orbslam3_cpp_fork_b200 import knnMatch2
    g = np.load(GOLDEN)
    qs = [g[f"{c}_q"] for c in CASES]; ts = [g[f"{c}_t"] for c in CASES]
    res = knnMatch2(qs, ts, ratio_test=True)                      # all five sets in ONE launch
    for c, (idx, dist, ratio) in zip(CASES, res):
        assert (idx == g[f"{c}_idx"]).all() and (dist == g[f"{c}_dist"]).all() and (ratio == g[f"{c}_ratio"]).all()
    idx, dist = knnMatch2([qs[0]], [ts[0]])[0]                    # single set, no ratio test
    assert (idx == g["random_idx"]).all() and (dist == g["random_dist"]).all()


def test_knn2_frame_sized_sets_vs_oracle():
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import knnMatch2
    rng = np.random.default_rng(4)
    qs, ts = [], []
    for s in range(6):                                            # six fisheye frames' lapping-area descriptors
        nq, nt = int(rng.integers(300, 1200)), int(rng.integers(300, 1200))
        t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
        q = t[rng.integers(0, nt, nq)] ^ (rng.integers(0, 256, (nq, 32), dtype=np.uint8) & rng.integers(0, 256, (nq, 32), dtype=np.uint8) & rng.integers(0, 256, (nq, 32), dtype=np.uint8))
        qs.append(q); ts.append(t)
    qs.append(np.zeros((0, 32), np.uint8)); ts.append(ts[0])       # empty query set
    qs.append(qs[0]); ts.append(np.zeros((0, 32), np.uint8))       # empty train set
    for (idx, dist, ratio), q, t in zip(knnMatch2(qs, ts, ratio_test=True), qs, ts):
        ri, rd = op.knn_match2(q, t)
        assert (idx == ri).all() and (dist == rd).all() and (ratio == op.fisheye_ratio_test(ri, rd)).all()


def test_distinctive_descriptors_vs_oracle():
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ComputeDistinctiveDescriptors
    obs = observations(2, n_points=300, max_obs=120) + [observations(3, 5, 1000)[4]]
    off = np.zeros(len(obs) + 1, np.int32); off[1:] = np.cumsum([len(o) for o in obs])
    ref = op.distinctive_descriptors(np.concatenate(obs), off)
    assert (ComputeDistinctiveDescriptors(obs) == ref).all()
