"""Experiment: does splitting one step's 256 frames over more handles (streams) improve overlap?
usage: python scripts/exp_streams.py  -> prints frames/s for 1, 2, 4 sub-batches."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from orbslam3_cpp_fork_b200 import ORBextractor, _lib
L_ = _lib.load()
B = 256
Lh, Rh = bench.make_pool(B)
Ld = torch.from_numpy(Lh).cuda(); Rd = torch.from_numpy(Rh).cuda()
for S in (1, 2, 4, 8):
    b = B // S
    hs = [(ORBextractor(bench.NFEAT, 1.2, 8, 20, 7, max_width=bench.W, max_height=bench.H, max_batch=b),
           ORBextractor(bench.NFEAT, 1.2, 8, 20, 7, max_width=bench.W, max_height=bench.H, max_batch=b)) for _ in range(S)]
    def step():
        for s, (eL, eR) in enumerate(hs):
            eL.extract_batch_async(Ld[s * b:(s + 1) * b]); eR.extract_batch_async(Rd[s * b:(s + 1) * b])
        for s, (eL, eR) in enumerate(hs):
            _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, bench.MB, bench.MBF))
    for _ in range(5): step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    N = 50
    for _ in range(N): step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"sub-batches {S} (batch {b} per handle): {B * N / dt:.0f} frames/s, {dt / N * 1e3:.3f} ms/step", flush=True)
    del hs
