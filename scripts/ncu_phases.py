#!/usr/bin/env python
"""Split a kernel's SASS (ncu --page source --csv) at BAR.SYNC and sum executed instructions,
stall samples and shared-memory wavefronts per region.  usage: ncu_phases.py rep [kernel-substr]"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
want = sys.argv[2] if len(sys.argv) > 2 else ""
rows = list(csv.reader(io.StringIO(out)))
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]; hdr = rows[i + 1]; j = i + 2
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            body.append(rows[j]); j += 1
        if want in name:
            print("==", name[:90])
            c = {k: hdr.index(k) for k in ["Source", "# Samples", "Instructions Executed", "L1 Wavefronts Shared", "L1 Wavefronts Shared Excessive", "Thread Instructions Executed"]}
            reg, acc, n = 0, [0, 0, 0, 0, 0], 0
            tot = sum(int(r[c["Instructions Executed"]] or 0) for r in body)
            def flush():
                print(f"  region {reg}: sass {n:4d}  inst {acc[0]:>12d} ({100*acc[0]/max(tot,1):5.1f}%)  samples {acc[1]:6d}  smem wf {acc[2]:>10d} excess {acc[3]:>10d}  thr/inst {acc[4]/max(acc[0],1):5.1f}")
            for r in body:
                acc[0] += int(r[c["Instructions Executed"]] or 0); acc[1] += int(r[c["# Samples"]] or 0)
                acc[2] += int(r[c["L1 Wavefronts Shared"]] or 0); acc[3] += int(r[c["L1 Wavefronts Shared Excessive"]] or 0)
                acc[4] += int(r[c["Thread Instructions Executed"]] or 0); n += 1
                if "BAR.SYNC" in r[c["Source"]] or "BAR.RED" in r[c["Source"]]:
                    flush(); reg += 1; acc = [0, 0, 0, 0, 0]; n = 0
            flush()
        i = j
    else:
        i += 1
