#!/bin/bash
# usage: scripts/ab.sh "ENV1=a ENV2=b" "ENV1=c" ...   -> one short bench line per environment setting
for cfg in "$@"; do
  env $cfg python bench.py --steps 100 --warmup 5 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$cfg', round(d['value']), round(d['e2e']['value']), round(d['ms_per_step'],4))"
done
