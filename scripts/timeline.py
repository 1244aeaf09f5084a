#!/usr/bin/env python
"""Kernel timeline of a few resident steps (torch.profiler = CUPTI activity records, process-wide, so the
kernels libcuorb launches on its own streams are included).  Prints one step's kernels in start order with
stream, start offset and duration, and the per-kernel totals; writes the raw list to gpurun_out/timeline.json."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from orbslam3_cpp_fork_b200 import ORBextractor, _lib

L_ = _lib.load()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
NSTEP = 6
Lh, Rh = bench.make_pool(B)
Ld, Rd = torch.from_numpy(Lh).cuda(), torch.from_numpy(Rh).cuda()
pairs = [(ORBextractor(bench.NFEAT, 1.2, 8, 20, 7, max_width=bench.W, max_height=bench.H, max_batch=B),
          ORBextractor(bench.NFEAT, 1.2, 8, 20, 7, max_width=bench.W, max_height=bench.H, max_batch=B)) for _ in range(2)]


def step(i):
    eL, eR = pairs[i & 1]
    eL.extract_batch_async(Ld); eR.extract_batch_async(Rd)
    _lib.check(L_.cuorb_stereo_match_batch_async(eL.handle, eR.handle, bench.MB, bench.MBF))


for i in range(4):
    step(i)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(NSTEP):
        step(i)
    torch.cuda.synchronize()
os.makedirs("gpurun_out", exist_ok=True)
prof.export_chrome_trace("gpurun_out/timeline_trace.json")
tr = json.load(open("gpurun_out/timeline_trace.json"))
ev = [e for e in tr["traceEvents"] if e.get("cat") == "kernel"]
ev.sort(key=lambda e: e["ts"])
t0 = ev[0]["ts"]
rows = [dict(name=e["name"].split("(")[0].replace("cuorb::", "")[:28], stream=e["args"].get("stream"), start=e["ts"] - t0, dur=e["dur"]) for e in ev]
json.dump(rows, open("gpurun_out/timeline.json", "w"))
os.remove("gpurun_out/timeline_trace.json")
total = rows[-1]["start"] + rows[-1]["dur"]
print(f"{NSTEP} steps, {len(rows)} kernels, span {total / 1e3:.3f} ms = {total / NSTEP / 1e3:.3f} ms/step")
# middle two steps
lo, hi = total * 2 / NSTEP, total * 4 / NSTEP
for r in rows:
    if lo <= r["start"] < hi:
        print(f"{r['start'] / 1e3:9.3f} ms  +{r['dur']:7.1f} us  s{r['stream']:<4} {r['name']}")
agg = {}
for r in rows:
    agg.setdefault(r["name"], [0, 0.0]); agg[r["name"]][0] += 1; agg[r["name"]][1] += r["dur"]
print("--- totals (sum of kernel durations, overlapping kernels both counted)")
for k, (n, d) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:30s} n={n:4d} sum={d / 1e3:8.3f} ms  per step {d / NSTEP / 1e3:7.3f} ms")
