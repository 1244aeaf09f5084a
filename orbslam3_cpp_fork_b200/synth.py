"""Seeded synthetic frames for parity tests and bench.py.

No datasets exist in the container, so inputs are procedural: band-limited noise texture plus a
few hundred random grey rectangles and checker patches (lots of FAST corners at every pyramid
level), then +-2 pixel noise.  Stereo pairs are the left image resampled through a smooth
disparity field d(x, y) in [2, 80] px, plus independent noise.  numpy only (PCG64 streams are
stable across numpy versions), so host tests, the CPU baseline and the GPU see identical bytes.
"""
from __future__ import annotations

import numpy as np


def _bilinear_up(lat: np.ndarray, h: int, w: int) -> np.ndarray:
    lh, lw = lat.shape
    ys = np.linspace(0, lh - 1.001, h, dtype=np.float32)
    xs = np.linspace(0, lw - 1.001, w, dtype=np.float32)
    y0 = ys.astype(np.int32); x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
    a = lat[y0][:, x0]; b = lat[y0][:, x0 + 1]; c = lat[y0 + 1][:, x0]; d = lat[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def synth_image(w: int, h: int, seed: int, n_shapes: int = 320, noise: int = 2) -> np.ndarray:
    """One u8 frame [h, w]."""
    rng = np.random.default_rng(0xC0FFEE + int(seed))
    img = np.full((h, w), 128.0, np.float32)
    for o, amp in enumerate((36.0, 24.0, 14.0, 8.0)):
        cell = 64 >> o
        lat = rng.standard_normal((h // cell + 3, w // cell + 3)).astype(np.float32)
        img += amp * _bilinear_up(lat, h, w)
    for _ in range(n_shapes):
        kind = rng.integers(0, 4)
        x0 = int(rng.integers(-10, w - 4)); y0 = int(rng.integers(-10, h - 4))
        sw = int(rng.integers(6, 70)); sh = int(rng.integers(6, 70))
        x1 = min(w, x0 + sw); y1 = min(h, y0 + sh); x0 = max(0, x0); y0 = max(0, y0)
        g = float(rng.integers(0, 256))
        c = int(rng.integers(4, 12))
        if x1 <= x0 or y1 <= y0:
            continue
        if kind < 3:
            img[y0:y1, x0:x1] = g
        else:  # checker patch
            yy, xx = np.mgrid[y0:y1, x0:x1]
            m = ((yy // c + xx // c) & 1).astype(bool)
            sub = img[y0:y1, x0:x1]
            sub[m] = g
            sub[~m] = 255.0 - g
    if noise:
        img += rng.integers(-noise, noise + 1, size=(h, w)).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def disparity_field(w: int, h: int, seed: int, dmin: float = 2.0, dmax: float = 80.0) -> np.ndarray:
    rng = np.random.default_rng(0xD15 + int(seed))
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    a, b = rng.uniform(-0.3, 0.3, 2)
    ph = rng.uniform(0, 6.28, 2)
    f = 0.5 + a * (xx / w - 0.5) + b * (yy / h - 0.5) \
        + 0.15 * np.sin(xx / w * 6.28 * 1.5 + ph[0]) * np.cos(yy / h * 6.28 + ph[1])
    f = np.clip(f, 0.0, 1.0)
    return (dmin + (dmax - dmin) * f).astype(np.float32)


def synth_stereo_pair(w: int, h: int, seed: int, noise: int = 2):
    """(left, right) u8 frames; right(x, y) = left(x + d(x, y), y), linear interpolation."""
    left = synth_image(w, h, seed, noise=noise)
    d = disparity_field(w, h, seed)
    xx = np.arange(w, dtype=np.float32)[None, :] + d
    x0 = np.floor(xx).astype(np.int32)
    fx = xx - x0
    x0c = np.clip(x0, 0, w - 1); x1c = np.clip(x0 + 1, 0, w - 1)
    rows = np.arange(h)[:, None]
    lf = left.astype(np.float32)
    right = lf[rows, x0c] * (1 - fx) + lf[rows, x1c] * fx
    rng = np.random.default_rng(0x51DE + int(seed))
    if noise:
        right += rng.integers(-noise, noise + 1, size=(h, w)).astype(np.float32)
    return left, np.clip(np.rint(right), 0, 255).astype(np.uint8)


def degenerate_image(w: int, h: int, kind: str, seed: int = 0) -> np.ndarray:
    """Edge-case frames: 'black' (no keypoints), 'flat' (few weak corners -> minThFAST retry),
    'noise' (far more candidates than the quota)."""
    rng = np.random.default_rng(0xBAD + int(seed))
    if kind == "black":
        return np.zeros((h, w), np.uint8)
    if kind == "flat":
        img = np.full((h, w), 120, np.uint8)
        for _ in range(12):  # faint squares: contrast 10 < iniThFAST, > minThFAST
            x0 = int(rng.integers(30, w - 60)); y0 = int(rng.integers(30, h - 60))
            img[y0:y0 + 20, x0:x0 + 20] = 131
        return img
    if kind == "noise":
        return rng.integers(0, 256, size=(h, w), dtype=np.uint8)
    raise ValueError(kind)


def color_frames(n: int, h: int, w: int, ch: int, seed: int) -> np.ndarray:
    """n seeded raw camera frames [n, h, w(, ch)] with structure in every channel (the gray synthetic
    frame plus per-channel offsets and noise) for the frame-preparation stage."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        g = synth_image(w, h, seed * 100 + i).astype(np.int32)
        if ch == 1:
            out.append(g.astype(np.uint8))
            continue
        chans = [np.clip(g + rng.integers(-30, 31, (h, w)) + 17 * (c - 1), 0, 255) for c in range(ch)]
        out.append(np.stack(chans, -1).astype(np.uint8))
    return np.stack(out)


def rectify_maps(w: int, h: int, seed: int = 0):
    """EuRoC-like stereo rectification maps (what cv::initUndistortRectifyMap would hand to cv::remap):
    radial distortion plus a small rotation, float32 [h, w] each."""
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    fx, fy, cx, cy = 0.61 * w, 0.95 * h, 0.49 * w, 0.52 * h
    x = (xx - cx) / fx; y = (yy - cy) / fy
    th = 0.012 + 0.001 * seed
    xr = x * np.cos(th) - y * np.sin(th) + 0.004; yr = x * np.sin(th) + y * np.cos(th) - 0.003
    r2 = xr * xr + yr * yr
    k = 1 - 0.2834 * r2 + 0.0739 * r2 * r2
    return (xr * k * fx + cx).astype(np.float32), (yr * k * fy + cy).astype(np.float32)
