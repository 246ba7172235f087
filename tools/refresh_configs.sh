#!/bin/bash
# usage (GPU box): tools/refresh_configs.sh -- GPU tests, then BASELINE configs[2] (1024 lights) and configs[3] (district, 4K) bench lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py --lights 1024 --no-stress --no-cpu-baseline > gpurun_out/bench_lights1024.json 2> gpurun_out/bench_lights1024.err; echo lights rc $?
timeout 900 python bench.py --workload district --no-cpu-baseline > gpurun_out/bench_district_4k.json 2> gpurun_out/bench_district_4k.err; echo district rc $?
python - <<'PY'
import json
for f in ("lights1024", "district_4k"):
    d=json.load(open(f"gpurun_out/bench_{f}.json")); print(f, round(d["value"],1), round(d["e2e"]["value"],1), d.get("blas"), d["clocks"]["reasons"])
PY
