#!/usr/bin/env python
"""A/B timing of cuorb_search_by_projection (cfg4: 50k MapPoints vs one frame) for a given libcuorb build.
usage: ab_projection.py [path/to/libcuorb.so] [M]"""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import numpy as np
    from orbslam3_cpp_fork_b200 import _lib
    if len(sys.argv) > 1 and sys.argv[1].endswith(".so"):
        _lib.LIB_PATH = os.path.abspath(sys.argv[1])
    M = int([a for a in sys.argv[1:] if a.isdigit()][0]) if any(a.isdigit() for a in sys.argv[1:]) else 50000
    import orbslam3_cpp_fork_b200 as cu
    from orbslam3_cpp_fork_b200.matcher import _c
    from tests.matching_cases import extracted_frame, mappoint_case
    L_ = _lib.load()
    e, k, d = extracted_frame()
    sf = e.tables()[0]
    c = mappoint_case(k, d, sf, M, 77 + M, with_uright=True)
    F = cu.Frame(k, d, sf, 752, 480, mvuRight=c["u_right"], mb=0.1, mbf=40.0)
    F.mvpMapPoints[c["bound"] > 0] = 10 ** 9
    F.mvpHasObs[c["bound"] > 0] = 1
    mps = cu.MapPointSoA(c["desc"], c["px"], c["py"], c["pxr"], c["level"], c["view_cos"], c["depth"],
                         (c["flags"] & 1) > 0, (c["flags"] & 2) > 0, (c["flags"] & 4) > 0)
    fvv, keep_alive = F.view()
    arrs = [_c(mps.descriptor, np.uint8), _c(mps.mTrackProjX, np.float32), _c(mps.mTrackProjY, np.float32),
            _c(mps.mTrackProjXR, np.float32), _c(mps.mnTrackScaleLevel, np.int32), _c(mps.mTrackViewCos, np.float32),
            _c(mps.mTrackDepth, np.float32), mps.flags()]
    best = np.zeros(M, np.int32)

    def call():
        return _lib.check(L_.cuorb_search_by_projection(0, C.byref(fvv), M, *[_lib.ptr(a) for a in arrs], 3.0, 0, 60.0, 0.8,
                                                        _lib.ptr(best)))
    for _ in range(20):
        n = call()
    ts = []
    for _ in range(200):
        t0 = time.perf_counter(); call(); ts.append(time.perf_counter() - t0)
    ts = np.sort(ts) * 1e3
    print(json.dumps({"lib": _lib.LIB_PATH, "M": M, "matches": int(n), "median_ms": float(ts[100]), "min_ms": float(ts[0]),
                      "p90_ms": float(ts[180]), "checksum": int(best.astype(np.int64).sum())}))


if __name__ == "__main__":
    main()
