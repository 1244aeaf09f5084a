"""GPU: libcuorb's UndistortKeyPoints / ComputeStereoFromRGBD (through the C ABI) — bit-exact against
the cv2 goldens and the CPU oracle."""
import numpy as np
import pytest

from tests.test_oracle_frameglue import CAMS, GOLDEN, keys_from

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cam", CAMS)
def test_undistort_matches_cv2_golden_and_oracle(cam):
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import UndistortKeyPoints
    g = np.load(GOLDEN)
    k = keys_from(g[f"{cam}_in"])
    K4, D = g[f"{cam}_K4"], g[f"{cam}_D"]
    out = UndistortKeyPoints(k, K4, D)
    assert (out["x"] == g[f"{cam}_out"][:, 0]).all() and (out["y"] == g[f"{cam}_out"][:, 1]).all()
    assert (out == op.undistort_keypoints(k, K4, D)).all()
    K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
    assert (UndistortKeyPoints(k, K, D) == out).all()           # 3x3 K form


def test_undistort_zero_k1_copy_and_empty():
    from orbslam3_cpp_fork_b200 import UndistortKeyPoints
    g = np.load(GOLDEN)
    k = keys_from(g["tum1_in"])
    assert (UndistortKeyPoints(k, g["tum1_K4"], [0.0, 0.3, 0.01, 0.0]) == k).all()
    assert len(UndistortKeyPoints(k[:0], g["tum1_K4"], g["tum1_D"])) == 0


def test_rgbd_stereo_host_and_device_depth():
    import torch
    from oracle import oracle_py as op
    from orbslam3_cpp_fork_b200 import ComputeStereoFromRGBD, UndistortKeyPoints
    rng = np.random.default_rng(5)
    w, h = 640, 480
    depth = rng.uniform(-0.2, 6.0, (h, w)).astype(np.float32)
    g = np.load(GOLDEN)
    k = keys_from(np.stack([rng.uniform(16, w - 16, 1200), rng.uniform(16, h - 16, 1200)], 1).astype(np.float32))
    ku = UndistortKeyPoints(k, g["tum1_K4"], g["tum1_D"])
    ru, rz = op.stereo_from_rgbd(k, ku, depth, 40.0)
    u, z = ComputeStereoFromRGBD(k, ku, depth, 40.0)
    assert (u == ru).all() and (z == rz).all()
    ud, zd = ComputeStereoFromRGBD(k, ku, torch.from_numpy(depth).cuda(), 40.0)
    assert (ud == ru).all() and (zd == rz).all()
    padded = np.zeros((h, w + 8), np.float32); padded[:, :w] = depth
    up, zp = ComputeStereoFromRGBD(k, ku, padded[:, :w], 40.0)
    assert (up == ru).all() and (zp == rz).all()


@pytest.mark.parametrize("seed,limit", [(1, 0.5), (2, 0.5), (3, 0.8)])
def test_is_in_frustum_matches_oracle_and_reference(seed, limit):
    """Frame::isInFrustum for 20k MapPoints: every output equal to the oracle and to the reference's own lifted
    body (oracle/_ref) bit for bit; differences, if any ever appear, can only sit at an FMA / logf bin edge and are
    counted separately here (tolerance of north_star: 1e-4 relative on floats, +-1 level at a PredictScale edge)."""
    from oracle import oracle_py as op
    from oracle import ref_py as rp
    from orbslam3_cpp_fork_b200 import _lib, isInFrustum
    from tests.matching_cases import frustum_case
    cam, P, N, min_d, max_d = frustum_case(20000, seed)
    ccam = _lib.FrustumCamera.from_buffer_copy(bytes(cam))
    g = isInFrustum(ccam, P, N, min_d, max_d, limit)
    names = dict(mbTrackInView="in_view", mTrackProjX="proj_x", mTrackProjY="proj_y", mTrackProjXR="proj_xr",
                 mnTrackScaleLevel="level", mTrackViewCos="view_cos", mTrackDepth="depth")
    checkers = [op.is_in_frustum(cam, P, N, min_d, max_d, limit)]
    if rp.available():
        checkers.append(rp.is_in_frustum(cam, P, N, min_d, max_d, limit))
    for o in checkers:
        edge = {k: int((g[k] != o[v]).sum()) for k, v in names.items()}
        assert g["count"] == o["count"], edge
        assert not any(edge.values()), f"bin-edge bucket (entries differing from the checker): {edge}"
    assert 2000 < g["count"] < 18000
