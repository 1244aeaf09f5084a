"""One call each of the matching entry points added late in round 2, for an `ncu --set full` capture:
cuorb_localmap_search (50k points), cuorb_search_by_projection (50k MapPoints), cuorb_search_by_projection_rig (20k).
usage: python scripts/profile_match.py   (under ncu --profile-from-start off: the capture window is opened here)"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import orbslam3_cpp_fork_b200 as cu
    from orbslam3_cpp_fork_b200 import _lib
    from tests.matching_cases import extracted_frame, localmap_case, mappoint_case
    from tests.rig_cases import W, H, rig_mappoints, rig_views
    e, k, d = extracted_frame()
    sf = np.asarray(e.tables()[0], np.float32)
    c = localmap_case(k, d, sf, 50000, 11)
    cam = _lib.FrustumCamera.from_buffer_copy(bytes(c["cam"]))
    m = cu.LocalMap(50000)
    m.update(0, c["P"], c["N"], c["min_d"], c["max_d"], c["desc"], c["bad"], c["has_obs"])

    def frame():
        F = cu.Frame(k, d, sf, 752, 480, mvuRight=c["u_right"], mb=0.1, mbf=c["mbf"])
        F.mvpMapPoints[c["bound"] > 0] = 10 ** 9; F.mvpHasObs[c["bound"] > 0] = 1
        return F
    mc = mappoint_case(k, d, sf, 50000, 77 + 50000, with_uright=True)
    mps = cu.MapPointSoA(mc["desc"], mc["px"], mc["py"], mc["pxr"], mc["level"], mc["view_cos"], mc["depth"],
                         (mc["flags"] & 1) > 0, (mc["flags"] & 2) > 0, (mc["flags"] & 4) > 0)
    rv = rig_views(9, nf=1200)
    q = rig_mappoints(rv, 20000, 1)
    args = [q[key] for key in ("desc", "px", "py", "level", "view_cos", "pxr", "pyr", "level_r", "view_cos_r", "depth", "flags")]

    def rig():
        L = cu.Frame(rv["k1"], rv["d1"], rv["sf"], W, H, mb=0.1, mbf=40.0); R = cu.Frame(rv["k2"], rv["d2"], rv["sf"], W, H, mb=0.1, mbf=40.0)
        return cu.RigFrame(L, R, rv["l2r"], rv["r2l"])
    M = cu.ORBmatcher(0.8, True)
    for _ in range(2):                                             # warm
        m.search(cam, frame(), c["skip_idx"], 0.5, 1.0, False, 50.0, 0.8, want_best=False, match_capacity=4096)
        M.SearchByProjection(frame(), mps, 3.0)
        M.SearchByProjectionRig(rig(), *args, 1.0, False, 40.0)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    n1 = m.search(cam, frame(), c["skip_idx"], 0.5, 1.0, False, 50.0, 0.8, want_best=False, match_capacity=4096)[0]
    n2 = M.SearchByProjection(frame(), mps, 3.0)[0]
    n3 = M.SearchByProjectionRig(rig(), *args, 1.0, False, 40.0)[0]
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    print("matches", n1, n2, n3)


if __name__ == "__main__":
    main()
