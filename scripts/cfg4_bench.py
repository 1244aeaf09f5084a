"""cfg4: SearchByProjection against 50k MapPoints through the C ABI, and the resident local map.
usage: python scripts/cfg4_bench.py out.json"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import bench_extras as bx
    r = bx.cfg4_projection(0, True)
    with open(sys.argv[1], "w") as f:
        json.dump(r, f, indent=1)
    print(json.dumps(r)[:3000])


if __name__ == "__main__":
    main()
