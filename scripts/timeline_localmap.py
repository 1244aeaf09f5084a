#!/usr/bin/env python
"""Device timeline of ONE cuorb_localmap_search call (50k MapPoints, 752x480 frame) and one cuorb_search_by_projection
call (cfg4) from CUPTI activity records (torch.profiler): start offset, duration, gap to the previous activity's end.
usage: python scripts/timeline_localmap.py > profiles/..._timeline_localmap.txt"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def dump(tr_path, calls, title):
    tr = json.load(open(tr_path)); os.remove(tr_path)
    ev = [x for x in tr["traceEvents"] if x.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    ev.sort(key=lambda x: x["ts"])
    n = len(ev) // calls
    ev = ev[(calls - 1) * n:]
    t0 = ev[0]["ts"]
    print(f"--- {title}: {len(ev)} device activities of one call (CUORB_PDL={os.environ.get('CUORB_PDL', '1 (default)')})")
    prev_end = None
    for x in ev:
        gap = 0.0 if prev_end is None else x["ts"] - prev_end
        prev_end = x["ts"] + x["dur"]
        name = x["name"].split("(")[0].replace("cuorb::", "").replace("void ", "")[:34]
        print(f"{(x['ts'] - t0):9.1f} us  +{x['dur']:6.1f} us  gap {gap:6.1f}  {name}")
    print(f"span {(ev[-1]['ts'] + ev[-1]['dur'] - t0):.1f} us")


def main():
    import numpy as np
    import torch
    from torch.profiler import profile, ProfilerActivity
    import orbslam3_cpp_fork_b200 as cu
    from orbslam3_cpp_fork_b200 import _lib
    from tests.matching_cases import extracted_frame, localmap_case, mappoint_case
    e, k, d = extracted_frame()
    sf = np.asarray(e.tables()[0], np.float32)
    M_ = int(os.environ.get("LM_POINTS", "50000"))
    c = localmap_case(k, d, sf, M_, 11)
    cam = _lib.FrustumCamera.from_buffer_copy(bytes(c["cam"]))
    m = cu.LocalMap(M_)
    m.update(0, c["P"], c["N"], c["min_d"], c["max_d"], c["desc"], c["bad"], c["has_obs"])

    def frame():
        F = cu.Frame(k, d, sf, 752, 480, mvuRight=c["u_right"], mb=0.1, mbf=c["mbf"])
        F.mvpMapPoints[c["bound"] > 0] = 10 ** 9; F.mvpHasObs[c["bound"] > 0] = 1
        return F
    mc = mappoint_case(k, d, sf, 50000, 77 + 50000, with_uright=True)
    mps = cu.MapPointSoA(mc["desc"], mc["px"], mc["py"], mc["pxr"], mc["level"], mc["view_cos"], mc["depth"],
                         (mc["flags"] & 1) > 0, (mc["flags"] & 2) > 0, (mc["flags"] & 4) > 0)
    M = cu.ORBmatcher(0.8, True)
    os.makedirs("gpurun_out", exist_ok=True)
    dk = torch.from_numpy(np.ascontiguousarray(k).view(np.uint8).copy()).cuda(); dd = torch.from_numpy(np.ascontiguousarray(d)).cuda()
    du = torch.from_numpy(np.ascontiguousarray(c["u_right"], np.float32)).cuda()
    torch.cuda.synchronize()
    for title, fn in ((f"cuorb_localmap_search_resident, {M_} points", lambda: m.search_resident(cam, frame(), dk.data_ptr(), dd.data_ptr(), du.data_ptr(), skip_idx=c["skip_idx"], thFarPoints=50.0, want_best=False, match_capacity=4096)),
                      (f"cuorb_localmap_search, {M_} points", lambda: m.search(cam, frame(), c["skip_idx"], 0.5, 1.0, False, 50.0, 0.8, want_best=False, match_capacity=4096)),
                      ("cuorb_search_by_projection, 50k MapPoints (cfg4)", lambda: M.SearchByProjection(frame(), mps, 3.0))):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
        path = "gpurun_out/_lm_trace.json"
        prof.export_chrome_trace(path)
        dump(path, 3, title)


if __name__ == "__main__":
    main()
