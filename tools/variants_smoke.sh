#!/bin/bash
# usage (GPU box): tools/variants_smoke.sh -- __graft_entry__.smoke(), then a short bench of the default library and every build/libprt_*.so
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
for f in prosper_b200/libprt_b200.so build/libprt_*.so; do
  PRT_LIB=$PWD/$f timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-stress 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); c=d['breakdown']['counters_per_step']; r=c['closestRays']+c['shadowRays']
print('$f', round(d['value'],1), d['breakdown']['stage_ms_last_step'], 'blas ms', round(d['tlas']['blas_build_ms_last'],2), 'nodes/ray', round(c['nodesVisited']/r,2), 'tris/ray', round(c['trisTested']/r,2))"
done
