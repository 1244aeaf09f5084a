"""Generate the golden vectors in tests/golden/ with the REAL OpenCV primitives (cv2 4.13.0) through
oracle/cv2_reference.py (the quadtree stage comes from the C++ oracle's std::list/std::sort code).
Run here (container with cv2); the .npz files are committed and travel to the GPU box.

    python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import cv2_reference as cr  # noqa: E402
from orbslam3_cpp_fork_b200 import synth  # noqa: E402

CASES = [  # name, w, h, nfeatures, seed, lapping, kind
    ("small_a", 320, 240, 500, 5, (0, 0), None),
    ("small_lap", 320, 240, 500, 6, (100, 200), None),
    ("small_flat", 320, 240, 300, 7, (0, 0), "flat"),
    ("euroc_mono", 752, 480, 1000, 2, (0, 1000), None),
    ("kitti", 1241, 376, 2000, 4, (0, 0), None),
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    import cv2
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name, w, h, nf, seed, lap, kind in CASES:
        img = synth.degenerate_image(w, h, kind, seed) if kind else synth.synth_image(w, h, seed)
        mono, k, d, st = cr.extract(img, nf, 1.2, 8, 20, 7, lap, return_stages=True)
        rec = dict(image=img, nfeatures=nf, lap=np.array(lap), mono=mono, kps=k, desc=d,
                   cv2_version=cv2.__version__,
                   pyr_sha=np.array([sha(p) for p in st["pyramid"]]),
                   blur_sha=np.array([sha(b) if b is not None else "" for b in st["blurred"]]),
                   cand_count=np.array([len(c) for c in st["candidates"]]),
                   cand_sha=np.array([sha(np.array(c, np.float32).reshape(-1, 3)) for c in st["candidates"]]))
        if w > 400:   # big inputs are regenerated from the seed; keep only their hash
            rec["image_sha"] = sha(img)
            rec["image_spec"] = np.array([w, h, seed])
            del rec["image"]
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **rec)
        print(name, "n", len(k), "mono", mono)


if __name__ == "__main__":
    main()
