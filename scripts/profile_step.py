#!/usr/bin/env python
"""One stereo step (2 x 256 extractions + 256 stereo matches) between cudaProfilerStart/Stop, for
`ncu --profile-from-start off --set full ...`: every kernel of the step is captured once."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from orbslam3_cpp_fork_b200 import ORBextractor, _lib

L_ = _lib.load()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
Lh, Rh = bench.make_pool(B)
Ld, Rd = torch.from_numpy(Lh).cuda(), torch.from_numpy(Rh).cuda()
eL = ORBextractor(bench.NFEAT, 1.2, 8, 20, 7, max_width=bench.W, max_height=bench.H, max_batch=B)
eR = ORBextractor(bench.NFEAT, 1.2, 8, 20, 7, max_width=bench.W, max_height=bench.H, max_batch=B)


def step():
    eL.extract_batch_async(Ld); eR.extract_batch_async(Rd)
    _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, bench.MB, bench.MBF))
    torch.cuda.synchronize()


for _ in range(3):
    step()
torch.cuda.cudart().cudaProfilerStart()
step()
torch.cuda.cudart().cudaProfilerStop()
print("profiled one step, batch", B)
