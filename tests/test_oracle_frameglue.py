"""CPU: oracle restatement of Frame::UndistortKeyPoints / ComputeStereoFromRGBD against cv2 goldens."""
import os

import numpy as np
import pytest

from oracle import oracle_py as op

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "frameglue_cv2.npz")
CAMS = ("euroc", "tum1", "barrel")


def keys_from(pts):
    k = np.zeros(len(pts), op.KP_DTYPE)
    k["x"] = pts[:, 0]; k["y"] = pts[:, 1]
    k["size"] = 31; k["angle"] = np.arange(len(pts)) % 360; k["response"] = 20; k["octave"] = np.arange(len(pts)) % 8
    k["class_id"] = -1
    return k


@pytest.mark.parametrize("cam", CAMS)
def test_oracle_undistort_matches_cv2_golden(cam):
    g = np.load(GOLDEN)
    k = keys_from(g[f"{cam}_in"])
    out = op.undistort_keypoints(k, g[f"{cam}_K4"], g[f"{cam}_D"])
    assert (out["x"] == g[f"{cam}_out"][:, 0]).all() and (out["y"] == g[f"{cam}_out"][:, 1]).all()
    for f in ("size", "angle", "response", "octave", "class_id"):
        assert (out[f] == k[f]).all()                       # only pt changes (src/Frame.cc:906-911)


def test_oracle_undistort_zero_k1_is_a_copy():
    g = np.load(GOLDEN)
    k = keys_from(g["euroc_in"])
    out = op.undistort_keypoints(k, g["euroc_K4"], np.array([0, 0.1, 0.01, 0.01], np.float32))
    assert (out == k).all()                                 # src/Frame.cc:887-890


def test_oracle_rgbd_semantics():
    rng = np.random.default_rng(3)
    h, w = 48, 64
    depth = rng.uniform(-0.5, 4.0, (h, w)).astype(np.float32)
    k = keys_from(np.stack([rng.uniform(0, w - 1, 300), rng.uniform(0, h - 1, 300)], 1).astype(np.float32))
    ku = k.copy(); ku["x"] += 0.37
    u, z = op.stereo_from_rgbd(k, ku, depth, 40.0)
    d = depth[k["y"].astype(np.int32), k["x"].astype(np.int32)]
    pos = d > 0
    assert (z[pos] == d[pos]).all() and (z[~pos] == -1).all() and (u[~pos] == -1).all()
    assert (u[pos] == (ku["x"][pos] - np.float32(40.0) / d[pos]).astype(np.float32)).all()
