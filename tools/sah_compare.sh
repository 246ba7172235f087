#!/bin/bash
# usage (GPU box): tools/sah_compare.sh  -- GPU tests of the binned-SAH builder, then the bench with each BLAS builder
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sah or ploc" 2>&1 | tail -5
for B in sah lbvh; do
  timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-stress --builder $B > gpurun_out/bench_builder_$B.json 2> gpurun_out/bench_builder_$B.err; echo $B rc $?
  python - <<PY
import json
d=json.load(open("gpurun_out/bench_builder_$B.json"))
c=d["breakdown"]["counters_per_step"]
print("$B", "Mrays/s", d["value"], d["breakdown"]["stage_ms_last_step"], "blas ms", d["tlas"].get("blas_build_ms_last"), "nodes/closest+shadow", c["nodesVisited"]/(c["closestRays"]+c["shadowRays"]), "tris", c["trisTested"]/(c["closestRays"]+c["shadowRays"]))
PY
done
