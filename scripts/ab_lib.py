#!/usr/bin/env python
"""A/B timing of a libcuorb build: per-stage CUDA-event times (one handle alone) and the resident two-pair loop of
bench.py's value leg, on device-generated cfg5 frames.  usage: ab_lib.py [path/to/libcuorb.so] [batch] [steps]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import numpy as np
    import torch
    from orbslam3_cpp_fork_b200 import _lib
    if len(sys.argv) > 1 and sys.argv[1].endswith(".so"):
        _lib.LIB_PATH = os.path.abspath(sys.argv[1])
    nums = [int(a) for a in sys.argv[1:] if a.isdigit()]
    B = nums[0] if nums else 256
    steps = nums[1] if len(nums) > 1 else 100
    from orbslam3_cpp_fork_b200 import ORBextractor, synth_cb
    L_ = _lib.load()
    W, H, MBF, FX = 752, 480, 47.90639384423901, 458.654
    pool = [synth_cb.stereo_batch_device(i * B, B, W, H) for i in range(8)]
    pairs = [(ORBextractor(1000, max_width=W, max_height=H, max_batch=B), ORBextractor(1000, max_width=W, max_height=H, max_batch=B))
             for _ in range(2)]

    def step(i):
        eL, eR = pairs[i & 1]
        l, r = pool[i % len(pool)]
        eL.extract_batch_async(l); eR.extract_batch_async(r)
        _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, MBF / FX, MBF))

    for i in range(4):
        step(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    for p in pairs:
        p[0].sync(); p[1].sync()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    eL, eR = pairs[0]
    eL.set_profiling(True)
    acc = {}
    for j in range(5):
        l, r = pool[j]
        eR.extract_batch_async(r); eR.sync()
        eL.extract_batch_async(l); eL.sync()
        _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, MBF / FX, MBF)); eL.sync()
        for k, v in eL.stage_ms().items():
            acc.setdefault(k, []).append(v)
    counts = eL.fetch()[0]
    print(json.dumps({"lib": _lib.LIB_PATH, "batch": B, "frames_per_s_resident": B * steps / dt, "ms_per_step": 1e3 * dt / steps,
                      "stage_ms": {k: round(float(np.median(v)), 4) for k, v in acc.items()}, "kp_checksum": int(counts.sum())}))


if __name__ == "__main__":
    main()
