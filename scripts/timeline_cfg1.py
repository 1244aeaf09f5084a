#!/usr/bin/env python
"""Kernel timeline of ONE synchronous single-frame extraction (cuorb_extract, 752x480, 1000 features) from CUPTI
activity records (torch.profiler): start offset, duration and the gap to the previous kernel's end, in stream order.
Run once as is and once with CUORB_PDL=0 to see what the dependent launches remove.
usage: python scripts/timeline_cfg1.py > profiles/..._timeline_cfg1.txt"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    from torch.profiler import profile, ProfilerActivity
    from orbslam3_cpp_fork_b200 import ORBextractor, synth
    img, _ = synth.synth_stereo_pair(752, 480, 4242)
    e = ORBextractor(1000, 1.2, 8, 20, 7, max_width=752, max_height=480, max_batch=1)
    for _ in range(5):
        e(img, None, (0, 1000))
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            e(img, None, (0, 1000))
        torch.cuda.synchronize()
    os.makedirs("gpurun_out", exist_ok=True)
    path = "gpurun_out/_cfg1_trace.json"
    prof.export_chrome_trace(path)
    tr = json.load(open(path)); os.remove(path)
    ev = [x for x in tr["traceEvents"] if x.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    ev.sort(key=lambda x: x["ts"])
    # the last of the three calls
    n = len(ev) // 3
    ev = ev[2 * n:]
    t0 = ev[0]["ts"]
    print(f"CUORB_PDL={os.environ.get('CUORB_PDL', '1 (default)')}: {len(ev)} device activities of one cuorb_extract call")
    prev_end = {}
    for x in ev:
        st = x["args"].get("stream")
        gap = x["ts"] - prev_end.get(st, x["ts"])
        prev_end[st] = x["ts"] + x["dur"]
        name = x["name"].split("(")[0].replace("cuorb::", "").replace("void ", "")[:34]
        print(f"{(x['ts'] - t0):9.1f} us  +{x['dur']:6.1f} us  gap {gap:6.1f}  s{st:<3} {name}")
    print(f"span {(ev[-1]['ts'] + ev[-1]['dur'] - t0):.1f} us")


if __name__ == "__main__":
    main()
