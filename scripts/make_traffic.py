#!/usr/bin/env python
"""profiles/traffic.json from an ncu launch list that carries dram__bytes_read/write.sum:
DRAM bytes per image and stage (pyramid = level0 + 7 resizes, fast, blur), averaged over the
launches found.  usage: make_traffic.py launches.csv batch out.json"""
import csv, json, sys, collections
src, batch, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
rows = [r for r in csv.reader(open(src)) if len(r) > 14 and r[0].isdigit()]
per = collections.defaultdict(lambda: collections.defaultdict(float))   # id -> metric -> value
name = {}
for r in rows:
    per[int(r[0])][r[12]] += float(r[14].replace(",", "")) * (1e6 if r[13] == "Mbyte" else 1e3 if r[13] == "Kbyte" else 1)
    name[int(r[0])] = r[4]
stage = {"k_pyr": "pyramid", "k_fast": "fast", "k_blur": "blur", "k_octree": "octree", "k_describe": "describe", "k_stereo_match": "stereo"}
tot = collections.defaultdict(float); cnt = collections.defaultdict(int)
for i, m in per.items():
    for k, s in stage.items():
        if k in name[i]:
            tot[s] += m.get("dram__bytes_read.sum", 0) + m.get("dram__bytes_write.sum", 0)
            cnt[s] += 1
# launches per extraction: pyramid 8, others 1
per_ext = {"pyramid": 8}
res = {}
for s in tot:
    n_ext = cnt[s] / per_ext.get(s, 1)
    res[s] = {"dram_bytes_per_image": tot[s] / n_ext / batch, "launches_seen": cnt[s], "source": src}
json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res, indent=1))
