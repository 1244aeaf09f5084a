"""Golden vectors for Frame::UndistortKeyPoints made with the REAL
cv2.undistortPoints (4.13.0), called the way the reference calls it (src/Frame.cc:903: K as both the
camera matrix and P, no R).  Run here (container with cv2):  python tests/golden/make_frameglue_golden.py"""
import os

import numpy as np

CAMERAS = {  # name -> (fx, fy, cx, cy), (k1, k2, p1, p2[, k3]), (w, h)   Examples/*/*.yaml of the reference
    "euroc": ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05), (752, 480)),
    "tum1": ((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314), (640, 480)),
    "barrel": ((300.0, 310.0, 320.0, 240.0), (-0.45, 0.32, 0.004, -0.003, -0.11), (640, 480)),
}


def main():
    import cv2
    rng = np.random.default_rng(20261105)
    rec = dict(cv2_version=cv2.__version__)
    for name, (K4, D, (w, h)) in CAMERAS.items():
        pts = np.stack([rng.uniform(0, w, 1500), rng.uniform(0, h, 1500)], 1).astype(np.float32)
        pts[:4] = [[0, 0], [w - 1, 0], [0, h - 1], [w - 1, h - 1]]
        K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
        Dm = np.array(D, np.float32)
        out = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, Dm, None, K).reshape(-1, 2)
        rec[f"{name}_K4"] = np.array(K4, np.float32); rec[f"{name}_D"] = Dm
        rec[f"{name}_in"] = pts; rec[f"{name}_out"] = out
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "frameglue_cv2.npz"), **rec)
    print("wrote frameglue_cv2.npz")


if __name__ == "__main__":
    main()
