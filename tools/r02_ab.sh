#!/bin/bash
# A/B of compile-time variants of libprt_b200.so.
#   here:        tools/r02_ab.sh build tag1="-DX=1" tag2="-DX=2 -DY" ...   -> prosper_b200/variants/libprt_<tag>.so
#   on the box:  tools/r02_ab.sh run [bench args]                         -> one line per variant (+ gpurun_out/ab_<tag>.json)
cd "$(dirname "$0")/.."
if [ "$1" = "build" ]; then
  shift; mkdir -p prosper_b200/variants; rm -f prosper_b200/variants/*.so
  for spec in "$@"; do
    tag=${spec%%=*}; flags=${spec#*=}
    ( cd prosper_b200/csrc && /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC $flags -shared -o ../variants/libprt_$tag.so prt_api.cu ) 2>&1 | grep -E "error" && echo "$tag FAILED"
    echo "built $tag ($flags)"
  done
  exit 0
fi
shift; mkdir -p gpurun_out
for lib in prosper_b200/variants/libprt_*.so; do
  tag=$(basename $lib .so); tag=${tag#libprt_}
  PRT_LIB=$PWD/$lib timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-stress --no-strong "$@" > gpurun_out/ab_$tag.json 2> gpurun_out/ab_$tag.err
  python - "$tag" <<PY
import json, sys
tag = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/ab_{tag}.json").read().strip().splitlines()[-1])
    print(tag, "Mrays/s %.1f" % d["value"], "ms/step %.1f" % d["ms_per_step"], d["breakdown"]["stage_ms_last_step"], "frac %.3f" % d["roofline"]["frac"], flush=True)
except Exception as e:
    print(tag, "FAILED", e, open(f"gpurun_out/ab_{tag}.err").read()[-600:], flush=True)
PY
done
