#!/usr/bin/env python
"""BASELINE cfg4 measurement helper: ORBmatcher::SearchByProjection of 50k local-map
MapPoints against one 752x480 / 1000-feature frame through the reference-facing
host API (host SoA buffers in, match indices out: H2D + kernels + D2H inside the timed call), with the CPU oracle
timed beside it on one core.  Run under `ncu -k regex:k_proj` to read the integer-pipe utilisation of the kernels.
usage: python scripts/bench_projection.py [M]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import orbslam3_cpp_fork_b200 as cu
from oracle import oracle_py as op
from tests.matching_cases import extracted_frame, mappoint_case

M = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
e, k, d = extracted_frame()
sf = e.tables()[0]
c = mappoint_case(k, d, sf, M, 77 + M, with_uright=True)
fv, keep = op.make_frame_view(k, d, c["u_right"], c["bound"], 752, 480, sf, mb=0.1, mbf=40.0)


def make_frame():
    F = cu.Frame(k, d, sf, 752, 480, mvuRight=c["u_right"], mb=0.1, mbf=40.0)
    F.mvpMapPoints[c["bound"] > 0] = 10 ** 9
    F.mvpHasObs[c["bound"] > 0] = 1
    return F


mps = cu.MapPointSoA(c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"], c["depth"],
                     (c["flags"] & 1) > 0, (c["flags"] & 2) > 0, (c["flags"] & 4) > 0)
m = cu.ORBmatcher(0.8)
n_gpu, best = m.SearchByProjection(make_frame(), mps, 3.0, False, 60.0)        # warm-up + parity
n_o, best_o = op.search_by_projection(fv, c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"],
                                      c["depth"], c["flags"], 3.0, False, 60.0, 0.8)
assert n_gpu == n_o and (best == best_o).all(), "parity"
frames = [make_frame() for _ in range(20)]
t0 = time.perf_counter()
for F in frames:
    m.SearchByProjection(F, mps, 3.0, False, 60.0)
dt_gpu = (time.perf_counter() - t0) / len(frames)
# the C-ABI call alone (what the C++ shim would time: host SoA in, best indices out)
import ctypes as C
from orbslam3_cpp_fork_b200 import _lib
from orbslam3_cpp_fork_b200.matcher import _c
L_ = _lib.load()
F = make_frame()
fvv, keep_alive = F.view()
flags = mps.flags()
arrs = [_c(mps.descriptor, np.uint8), _c(mps.mTrackProjX, np.float32), _c(mps.mTrackProjY, np.float32),
        _c(mps.mTrackProjXR, np.float32), _c(mps.mnTrackScaleLevel, np.int32), _c(mps.mTrackViewCos, np.float32),
        _c(mps.mTrackDepth, np.float32), flags]
best_c = np.zeros(M, np.int32)
def c_call():
    return _lib.check(L_.cuorb_search_by_projection(0, C.byref(fvv), M, *[_lib.ptr(a) for a in arrs], 3.0, 0, 60.0, 0.8,
                                                    _lib.ptr(best_c)))
c_call()
assert (best_c == best_o).all()
t0 = time.perf_counter()
for _ in range(50):
    rounds = c_call()
dt_c = (time.perf_counter() - t0) / 50
t0 = time.perf_counter()
for _ in range(5):
    op.search_by_projection(fv, c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"], c["depth"], c["flags"],
                            3.0, False, 60.0, 0.8)
dt_cpu = (time.perf_counter() - t0) / 5
out = {"workload": f"cfg4: SearchByProjection(F, {M} MapPoints), frame 752x480 / {len(k)} keypoints, th 3",
       "matches": int(n_gpu), "parity": "match count and every best index equal to the oracle",
       "cuorb_c_abi_ms_per_call_host_buffers": dt_c * 1e3, "mappoints_per_s": M / dt_c,
       "python_wrapper_ms_per_call": dt_gpu * 1e3,
       "cpu_oracle_ms_per_call_1core": dt_cpu * 1e3, "cpu_mappoints_per_s_1core": M / dt_cpu}

print(json.dumps(out))
