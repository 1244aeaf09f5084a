#!/bin/bash
# usage: scripts/ab.sh "ENV1=a ENV2=b" "ENV1=c" ...   -> one short bench line per environment setting
# (resident frames/s, e2e frames/s, ms per step, serialised stage times per 256 images)
for cfg in "$@"; do
  env $cfg python bench.py --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); st=d['roofline']['stage_ms_per_step']
print('$cfg', round(d['value']), round(d['e2e']['value']), round(d['ms_per_step'],4), {k:round(v,3) for k,v in st.items()})"
done
