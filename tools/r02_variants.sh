#!/bin/bash
# usage (on the GPU box): tools/r02_variants.sh [notests] -- GPU parity tests, then the bench under the pooled-trace experiment switches
mkdir -p gpurun_out
[ "$1" = "notests" ] || (timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8) | tee gpurun_out/r02_gpu_tests.log
run() { # tag, env...
  local tag=$1; shift
  local envs=() extra=()
  while [ $# -gt 0 ]; do if [ "$1" = "--" ]; then shift; extra=("$@"); break; fi; envs+=("$1"); shift; done
  env "${envs[@]}" timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-stress "${extra[@]}" > gpurun_out/v_$tag.json 2> gpurun_out/v_$tag.err
  python - "$tag" <<PY
import json, sys
tag = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/v_{tag}.json"))
    print(tag, "Mrays/s %.1f" % d["value"], "ms/step %.1f" % d["ms_per_step"], d["breakdown"]["stage_ms_last_step"], "frac %.3f" % d["roofline"]["frac"], flush=True)
except Exception as e:
    print(tag, "FAILED", e, open(f"gpurun_out/v_{tag}.err").read()[-600:], flush=True)
PY
}
if [ -f tools/r02_variant_list.sh ]; then source tools/r02_variant_list.sh; fi
