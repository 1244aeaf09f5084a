"""Counter-based synthetic stereo frames (SURVEY.md §8d "cfg5"): every pixel is a pure integer function of
(frame id, x, y), so the same bytes come out of this numpy statement and of the CUDA generator
(csrc/synth.cu, cuorb_synth_stereo) — which is what lets bench.py run the 100k-pair sweep over DISTINCT
frames produced on the device, with a host-certified subset (tests/test_gpu_synth.py).

Integer-only on purpose (no float rounding to disagree about):
  hash     splitmix64(seed * K1 + stream * K2 + index)
  texture  4 octaves of value noise on lattices of 64/32/16/8 px, bilinear in 8.8 fixed point
  shapes   320 rectangles / checker patches, painter's order (the last shape covering a pixel wins)
  noise    +-2 grey levels per pixel, independent per eye
  right    left (before noise) resampled along x through a disparity field in 1/256 px,
           d = clamp(plane + 256-px value noise, 2 px, 80 px), linear interpolation, rounded
"""
from __future__ import annotations

import numpy as np

K1 = np.uint64(0x9E3779B97F4A7C15)
K2 = np.uint64(0xD1B54A32D192ED03)
SEED0 = 0xC0FFEE
N_SHAPES = 320
AMPS = (36, 24, 14, 8)          # per octave, cell = 64 >> octave
S_LATTICE, S_SHAPE, S_NOISE_L, S_NOISE_R, S_DISP = 1, 16, 7, 8, 9


def splitmix64(x):
    with np.errstate(over="ignore"):
        z = (np.asarray(x, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def rnd(seed, stream, idx):
    with np.errstate(over="ignore"):
        return splitmix64(np.uint64(seed) * K1 + np.uint64(stream) * K2 + np.asarray(idx, dtype=np.uint64))


def _value_noise(seed, stream, cell_log2, xs, ys):
    """8.8 fixed-point bilinear value noise, lattice values in [-128, 127]; xs / ys int64 arrays (broadcastable)."""
    cell = 1 << cell_log2
    ix, iy = xs >> cell_log2, ys >> cell_log2
    fx = (xs & (cell - 1)) << (8 - cell_log2)
    fy = (ys & (cell - 1)) << (8 - cell_log2)

    def lat(i, j):
        return (rnd(seed, stream, (j.astype(np.uint64) << np.uint64(16)) + i.astype(np.uint64)) >> np.uint64(56)).astype(np.int64) - 128
    a, b, c, d = lat(ix, iy), lat(ix + 1, iy), lat(ix, iy + 1), lat(ix + 1, iy + 1)
    top = a * (256 - fx) + b * fx
    bot = c * (256 - fx) + d * fx
    return (top * (256 - fy) + bot * fy) >> 16


def shape_params(seed, w, h):
    """[N_SHAPES, 7] int64: kind, x0, y0, x1, y1 (clipped, exclusive), grey, checker size."""
    k = np.arange(N_SHAPES, dtype=np.uint64)

    def f(field, lo, n):
        return lo + (rnd(seed, S_SHAPE + field, k) % np.uint64(n)).astype(np.int64)
    kind = f(0, 0, 4)
    x0 = f(1, -10, w - 4 + 10); y0 = f(2, -10, h - 4 + 10)
    sw = f(3, 6, 64); sh = f(4, 6, 64)
    g = f(5, 0, 256); c = f(6, 4, 8)
    x1 = np.minimum(w, x0 + sw); y1 = np.minimum(h, y0 + sh)
    x0 = np.maximum(0, x0); y0 = np.maximum(0, y0)
    return np.stack([kind, x0, y0, x1, y1, g, c], axis=1)


def clean_left(frame_id, w, h):
    """left image before pixel noise, int64 [h, w] in 0..255"""
    seed = SEED0 + int(frame_id)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.int64)
    img = np.full((h, w), 128, np.int64)
    for o, amp in enumerate(AMPS):
        img += (amp * _value_noise(seed, S_LATTICE + o, 6 - o, xs, ys)) >> 7
    img = np.clip(img, 0, 255)
    for kind, x0, y0, x1, y1, g, c in shape_params(seed, w, h):
        if x1 <= x0 or y1 <= y0:
            continue
        if kind < 3:
            img[y0:y1, x0:x1] = g
        else:
            yy, xx = np.mgrid[y0:y1, x0:x1]
            m = ((yy // c + xx // c) & 1).astype(bool)
            img[y0:y1, x0:x1] = np.where(m, g, 255 - g)
    return img


def _noise(seed, stream, w, h):
    idx = np.arange(w * h, dtype=np.uint64).reshape(h, w)
    return (rnd(seed, stream, idx) % np.uint64(5)).astype(np.int64) - 2


def disparity256(frame_id, w, h):
    """disparity in 1/256 px, int64 [h, w], clamped to [2, 80] px"""
    seed = SEED0 + int(frame_id)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.int64)
    r = rnd(seed, S_DISP, np.arange(3, dtype=np.uint64))
    base = 20 * 256 + int(r[0] % np.uint64(30 * 256))            # 20 .. 50 px
    ax = int(r[1] % np.uint64(2049)) - 1024                      # slope: +-4 px across 256 px, in 1/256 px per px
    ay = int(r[2] % np.uint64(2049)) - 1024
    d = base + ((ax * xs) >> 8) + ((ay * ys) >> 8) + 16 * _value_noise(seed, S_DISP + 1, 8, xs, ys)
    return np.clip(d, 2 * 256, 80 * 256)


def stereo_pair(frame_id, w=752, h=480):
    """(left, right) u8 [h, w] of frame `frame_id`."""
    seed = SEED0 + int(frame_id)
    clean = clean_left(frame_id, w, h)
    left = np.clip(clean + _noise(seed, S_NOISE_L, w, h), 0, 255).astype(np.uint8)
    d = disparity256(frame_id, w, h)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.int64)
    xi = xs + (d >> 8)
    fr = d & 255
    a = clean[ys, np.minimum(xi, w - 1)]
    b = clean[ys, np.minimum(xi + 1, w - 1)]
    res = (a * (256 - fr) + b * fr + 128) >> 8
    right = np.clip(res + _noise(seed, S_NOISE_R, w, h), 0, 255).astype(np.uint8)
    return left, right


def stereo_batch(first_frame_id, n, w=752, h=480):
    L = np.empty((n, h, w), np.uint8); R = np.empty((n, h, w), np.uint8)
    for i in range(n):
        L[i], R[i] = stereo_pair(first_frame_id + i, w, h)
    return L, R


def stereo_batch_device(first_frame_id, n, w=752, h=480, device=0, out=None):
    """The same frames produced by the CUDA generator (cuorb_synth_stereo): (left, right) torch uint8 CUDA
    tensors [n, h, w].  `out` = (left, right, scratch) tensors to fill in place."""
    import torch
    from . import _lib
    L_ = _lib.load()
    dev = torch.device("cuda", device)
    if out is None:
        out = tuple(torch.empty((n, h, w), dtype=torch.uint8, device=dev) for _ in range(3))
    left, right, scratch = out
    st = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(L_.cuorb_synth_stereo(device, int(first_frame_id), n, w, h, _lib.ptr(left.data_ptr()),
                                     _lib.ptr(right.data_ptr()), _lib.ptr(scratch.data_ptr()), _lib.ptr(st)))
    return left, right
