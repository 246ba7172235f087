#!/bin/bash
# usage (GPU box): tools/sah_variants2.sh -- short bench of every build/libprt_*.so, then the district workload
mkdir -p gpurun_out
for f in build/libprt_*.so; do
  PRT_LIB=$PWD/$f timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-stress 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); c=d['breakdown']['counters_per_step']; r=c['closestRays']+c['shadowRays']
print('$f', round(d['value'],1), d['breakdown']['stage_ms_last_step'], 'blas ms', round(d['tlas']['blas_build_ms_last'],2), 'nodes/ray', round(c['nodesVisited']/r,2), 'tris/ray', round(c['trisTested']/r,2))"
done
timeout 900 python bench.py --workload district --no-cpu-baseline > gpurun_out/bench_district.json 2> gpurun_out/bench_district.err; echo district rc $?
tail -c 1500 gpurun_out/bench_district.json; tail -3 gpurun_out/bench_district.err
