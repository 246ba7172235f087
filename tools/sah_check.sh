#!/bin/bash
# usage (GPU box): tools/sah_check.sh -- GPU tests, a short atrium bench and the district workload with the default builder
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-stress 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); c=d['breakdown']['counters_per_step']; r=c['closestRays']+c['shadowRays']
print('atrium', round(d['value'],1), d['breakdown']['stage_ms_last_step'], 'blas ms', round(d['tlas']['blas_build_ms_last'],2), 'nodes/ray', round(c['nodesVisited']/r,2), 'tris/ray', round(c['trisTested']/r,2))"
timeout 900 python bench.py --workload district --no-cpu-baseline > gpurun_out/bench_district.json 2> gpurun_out/bench_district.err; echo district rc $?
python -c "
import json; d=json.load(open('gpurun_out/bench_district.json')); print('district', d['value'], d['e2e']['value'], d['tlas'], d['blas'])"
