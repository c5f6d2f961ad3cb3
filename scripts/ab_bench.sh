#!/bin/bash
# A/B runs of the one-GPU bench under environment overrides (no rebuild).  usage: ab_bench.sh <out.log> "ENV=.. ENV=.." ...
out=$1; shift
mkdir -p gpurun_out
: > "$out"
for cfg in "$@"; do
  echo "== $cfg" >> "$out"
  env $cfg timeout 300 python bench.py --steps 40 --warmup 8 2>>"$out.err" | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['ms_per_step'], d['config'].get('step_ms_min_med_max'), 'e2e', d['e2e']['ms_per_step'], 'launches', d.get('launches_per_step'))
" >> "$out"
done
