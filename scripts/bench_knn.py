"""N4 measurement helper: 64 fisheye-frame-sized brute-force 2-NN problems (1000 x 1000 descriptors each)
in one k_knn2 launch through the C ABI.  Run under ncu to read the integer-pipe utilisation."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from orbslam3_cpp_fork_b200 import knnMatch2
rng = np.random.default_rng(0)
qs = [rng.integers(0, 256, (1000, 32), dtype=np.uint8) for _ in range(64)]
ts = [rng.integers(0, 256, (1000, 32), dtype=np.uint8) for _ in range(64)]
knnMatch2(qs, ts, True)
t0 = time.perf_counter()
for _ in range(5):
    knnMatch2(qs, ts, True)
dt = (time.perf_counter() - t0) / 5
print(f"64 x (1000 x 1000) pairs incl. host<->device copies: {dt * 1e3:.3f} ms/call, {64e6 / dt / 1e9:.2f} G pairs/s")
