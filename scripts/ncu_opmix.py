#!/usr/bin/env python
"""Executed-instruction mix by SASS opcode for one kernel of a .ncu-rep (source page).
usage: ncu_opmix.py rep kernel-substr [nth]"""
import csv, io, subprocess, sys, collections
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
want = sys.argv[2]; nth = int(sys.argv[3]) if len(sys.argv) > 3 else 0
rows = list(csv.reader(io.StringIO(out)))
i = 0; seen = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]; hdr = rows[i + 1]; j = i + 2; body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            body.append(rows[j]); j += 1
        if want in name:
            if seen == nth:
                ci, cs = hdr.index("Instructions Executed"), hdr.index("Source")
                mix = collections.Counter(); tot = 0
                for r in body:
                    op = r[cs].split()[0] if not r[cs].strip().startswith("@") else r[cs].split()[1]
                    op = op.split(".")[0]; n = int(r[ci] or 0); mix[op] += n; tot += n
                print(name[:80], "total", tot)
                for op, n in mix.most_common(22): print(f"  {op:12s} {n:>12d} {100*n/tot:5.1f}%")
                break
            seen += 1
        i = j
    else: i += 1
