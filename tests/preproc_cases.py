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


from orbslam3_cpp_fork_b200.synth import color_frames, rectify_maps  # noqa: E402,F401
