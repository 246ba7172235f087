#!/bin/bash
# usage (GPU box): tools/variants.sh  -- a short bench of every build/libprt_*.so (kernel experiments)
for f in build/libprt_*.so; do
  PRT_LIB=$PWD/$f timeout 600 python bench.py --steps 1 --warmup 2 --no-cpu-baseline --no-stress 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$f', round(d['value'],1), d['breakdown']['stage_ms_last_step'])"
done
