#!/usr/bin/env python
"""Top CUDA source lines of one kernel by stall samples (ncu --page source --print-source cuda --csv; needs -lineinfo and
--import-source on).  usage: ncu_lines.py rep kernel-substr [top-n]"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--print-source", "cuda", "--csv"], capture_output=True, text=True).stdout
want = sys.argv[2]; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
rows = list(csv.reader(io.StringIO(out)))
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]; hdr = rows[i + 1]; j = i + 2
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            body.append(rows[j]); j += 1
        if want in name and "# Samples" in hdr:
            cs, ci, cl = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Source")
            ln = hdr.index("#") if "#" in hdr else None
            tot = sum(int(r[cs] or 0) for r in body if len(r) > cs)
            print("==", name[:100], "total samples", tot)
            ranked = sorted((r for r in body if len(r) > cs and (r[cs] or "0") != "0"), key=lambda r: -int(r[cs] or 0))[:top]
            for r in ranked:
                print(f"  {100 * int(r[cs]) / max(tot, 1):5.1f}%  inst {int(r[ci] or 0):>9d}  L{r[ln] if ln is not None else '?':>4}  {r[cl].strip()[:150]}")
        i = j
    else:
        i += 1
