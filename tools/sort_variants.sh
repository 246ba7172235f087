#!/bin/bash
# usage (GPU box): tools/sort_variants.sh "0 1 3 7"  -- bench with each PRT_RAY_SORT mode (ray reordering experiment)
for m in ${1:-0 1 3 7}; do
  PRT_RAY_SORT=$m timeout 600 python bench.py --steps 1 --warmup 2 --no-cpu-baseline --no-stress 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); b=d['breakdown']; print('sort=$m', round(d['value'],1), {k: round(v,1) for k,v in b['stage_ms_last_step'].items()}, 'launches', d['gpu_launches'])"
done
