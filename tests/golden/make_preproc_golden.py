"""Golden vectors for the pre-extraction image ops, made with the REAL OpenCV
(cv2 4.13.0): cv2.remap(INTER_LINEAR) with cv2.initUndistortRectifyMap maps, cv2.resize, cv2.cvtColor.
Run here (container with cv2); the .npz files are committed and travel to the GPU box.

    python tests/golden/make_preproc_golden.py
"""
import hashlib
import os

import numpy as np


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def frame(rng, h, w, ch):
    yy, xx = np.mgrid[0:h, 0:w]
    base = (96 + 60 * np.sin(xx / 9.0) * np.cos(yy / 7.0)).astype(np.float32)
    img = np.stack([base + rng.integers(-40, 41, (h, w)) for _ in range(ch)], -1)
    img[h // 4:h // 2, w // 3:w // 2] = rng.integers(0, 256, ch)
    img = np.clip(img, 0, 255).astype(np.uint8)
    return img[..., 0].copy() if ch == 1 else img


def main():
    import cv2
    out_dir = os.path.dirname(os.path.abspath(__file__))
    rng = np.random.default_rng(20261104)
    sw, sh = 126, 80                       # EuRoC geometry / 6
    K = np.array([[458.654 / 6, 0, 367.215 / 6], [0, 457.296 / 6, 248.375 / 6], [0, 0, 1]])
    D = np.array([-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0])
    R = cv2.Rodrigues(np.array([0.012, -0.021, 0.006]))[0]
    P = np.array([[435.2 / 6, 0, 367.4 / 6, 0], [0, 435.2 / 6, 252.2 / 6, 0], [0, 0, 1, 0]])
    m1, m2 = cv2.initUndistortRectifyMap(K, D, R, P, (sw, sh), cv2.CV_32F)
    # a map that leaves the source on every side (BORDER_CONSTANT taps) and has negative coordinates
    yy, xx = np.mgrid[0:64, 0:100].astype(np.float32)
    w1 = (xx * 1.45 - 14.3 + 2.5 * np.sin(yy / 11)).astype(np.float32)
    w2 = (yy * 1.42 - 9.6 + 1.5 * np.cos(xx / 13)).astype(np.float32)
    codes = {1: cv2.COLOR_RGB2GRAY, 2: cv2.COLOR_BGR2GRAY, 3: cv2.COLOR_RGBA2GRAY, 4: cv2.COLOR_BGRA2GRAY}
    rec = dict(cv2_version=cv2.__version__, rmap_x=m1, rmap_y=m2, wmap_x=w1, wmap_y=w2)
    for fmt, ch in ((0, 1), (1, 3), (2, 3), (3, 4), (4, 4)):
        src = frame(rng, sh, sw, ch)
        rec[f"src{fmt}"] = src
        gray = (lambda a: a) if fmt == 0 else (lambda a, c=codes[fmt]: cv2.cvtColor(a, c))
        outs = dict(none=gray(src), remap=gray(cv2.remap(src, m1, m2, cv2.INTER_LINEAR)),
                    wild=gray(cv2.remap(src, w1, w2, cv2.INTER_LINEAR)),
                    down=gray(cv2.resize(src, (95, 60))),      # arbitrary ratio
                    half=gray(cv2.resize(src, (63, 40))),      # exact 2x2 -> INTER_AREA reroute
                    up=gray(cv2.resize(src, (154, 100))))      # upscaling
        for k, v in outs.items():
            rec[f"{k}{fmt}_shape"] = np.array(v.shape)
            rec[f"{k}{fmt}_sha"] = sha(v)
            if fmt == 2:                                       # full arrays for one format (diagnosable)
                rec[f"{k}{fmt}"] = v
    np.savez_compressed(os.path.join(out_dir, "preproc_cv2.npz"), **rec)
    print("wrote preproc_cv2.npz", len(rec), "entries")


if __name__ == "__main__":
    main()
