"""Local-map call (cuorb_localmap_search, host frame and device-resident frame) at the map sizes a tracking thread
really holds (2k - 10k points): ms per call next to the oracle on one core.
usage: python scripts/localmap_small.py"""
import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench_extras as bx
for M in (2000, 4000, 10000, 50000):
    r = bx.localmap_track(0, with_cpu=True, M=M)
    print(M, r["points_in_view"], r["matches"], r["ms_per_call"]["median_ms"], r["frame_left_on_device"]["ms_per_call"]["median_ms"], r["cpu_port_ms_per_call_1core"], r["equal_to_cpu"])
