#!/usr/bin/env python
"""One stereo step (2 x B extractions + B stereo matches) between cudaProfilerStart/Stop, for
`ncu --profile-from-start off --set full ...`: every kernel of the step is captured once.
usage: profile_step.py [cfg5|cfg1|cfg3] [batch]
  cfg5  752x480 / 1000 features (bench.py's headline workload; default batch 256)
  cfg1  752x480 / 1000 features, batch 1 (the single-frame call of a live SLAM run)
  cfg3  1241x376 / 2000 features (KITTI shape; default batch 96)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from orbslam3_cpp_fork_b200 import ORBextractor, _lib, synth_cb
    cfg = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].isdigit() else "cfg5"
    rest = [a for a in sys.argv[1:] if a.isdigit()]
    W, H, NF, B, FX, MBF = {"cfg5": (752, 480, 1000, 256, 458.654, 47.90639384423901),
                            "cfg1": (752, 480, 1000, 1, 458.654, 47.90639384423901),
                            "cfg3": (1241, 376, 2000, 96, 718.856, 386.1448)}[cfg]
    if rest:
        B = int(rest[0])
    L_ = _lib.load()
    Ld, Rd = synth_cb.stereo_batch_device(0, B, W, H)
    eL = ORBextractor(NF, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)
    eR = ORBextractor(NF, 1.2, 8, 20, 7, max_width=W, max_height=H, max_batch=B)

    def step():
        eL.extract_batch_async(Ld); eR.extract_batch_async(Rd)
        _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, MBF / FX, MBF))
        torch.cuda.synchronize()

    for _ in range(3):
        step()
    torch.cuda.cudart().cudaProfilerStart()
    step()
    torch.cuda.cudart().cudaProfilerStop()
    print("profiled one step:", cfg, "batch", B)


if __name__ == "__main__":
    main()
