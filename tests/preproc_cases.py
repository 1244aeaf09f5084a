"""Shared case table for the pre-extraction image-op tests."""
import hashlib
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "preproc_cv2.npz")
FMT_CH = {0: 1, 1: 3, 2: 3, 3: 4, 4: 4}
# name -> (mode, dsize or map keys)
CASES = {
    "none": (0, None), "remap": (1, ("rmap_x", "rmap_y")), "wild": (1, ("wmap_x", "wmap_y")),
    "down": (2, (95, 60)), "half": (2, (63, 40)), "up": (2, (154, 100)),
}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def color_frames(n, h, w, ch, seed):
    """n seeded frames [n, h, w(, ch)] with structure in every channel (not plain noise)."""
    from orbslam3_cpp_fork_b200 import synth
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        g = synth.synth_image(w, h, seed * 100 + i).astype(np.int32)
        if ch == 1:
            out.append(g.astype(np.uint8))
            continue
        chans = [np.clip(g + rng.integers(-30, 31, (h, w)) + 17 * (c - 1), 0, 255) for c in range(ch)]
        out.append(np.stack(chans, -1).astype(np.uint8))
    return np.stack(out)


def rectify_maps(w, h, seed=0):
    """EuRoC-like rectification maps without cv2: radial distortion + a small rotation, float32."""
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    fx, fy, cx, cy = 0.61 * w, 0.95 * h, 0.49 * w, 0.52 * h
    x = (xx - cx) / fx; y = (yy - cy) / fy
    th = 0.012 + 0.001 * seed
    xr = x * np.cos(th) - y * np.sin(th) + 0.004; yr = x * np.sin(th) + y * np.cos(th) - 0.003
    r2 = xr * xr + yr * yr
    k = 1 - 0.2834 * r2 + 0.0739 * r2 * r2
    return (xr * k * fx + cx).astype(np.float32), (yr * k * fy + cy).astype(np.float32)
