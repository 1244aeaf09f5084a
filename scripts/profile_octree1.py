"""One synchronous single-frame extraction inside a profiler window (for ncu -k regex:k_octree)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

def main():
    import torch
    from orbslam3_cpp_fork_b200 import ORBextractor, synth
    img, _ = synth.synth_stereo_pair(752, 480, 4242)
    e = ORBextractor(1000, 1.2, 8, 20, 7, max_width=752, max_height=480, max_batch=1)
    for _ in range(3):
        e(img, None, (0, 1000))
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    e(img, None, (0, 1000))
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()

if __name__ == "__main__":
    main()
