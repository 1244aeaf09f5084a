"""cfg1: single-frame latency of the synchronous C-ABI calls, wall clock, on cuda:0.
usage: python scripts/cfg1_latency.py out.json   (CUORB_PDL=0 turns the dependent launches off for an A/B)"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import bench_extras as bx
    r = bx.cfg1_latency(0, False)
    r["CUORB_PDL"] = os.environ.get("CUORB_PDL", "1 (default)")
    with open(sys.argv[1], "w") as f:
        json.dump(r, f, indent=1)
    print(json.dumps({k: (v["median_ms"] if isinstance(v, dict) and "median_ms" in v else v) for k, v in r.items()
                      if k != "workload"}))


if __name__ == "__main__":
    main()
