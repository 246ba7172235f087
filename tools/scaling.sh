#!/bin/bash
# usage (GPU box, gpurun --gpus 8): tools/scaling.sh  -- bench.py at 1, 2, 4 and 8 GPUs back to back, one JSON line each
mkdir -p gpurun_out
F="--steps 3 --warmup 3 --no-cpu-baseline --no-stress --no-roofline"
timeout 600 python bench.py --gpus 1 $F > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; echo n1 rc $?
for N in 2 4 8; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N $F > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo n$N rc $?
done
python - <<'PY'
import json
out=[]
for n in (1,2,4,8):
    try:
        d=json.loads(open(f"gpurun_out/scale_n{n}.json").read().strip().splitlines()[-1])
    except Exception as e:
        print(n, "missing", e); continue
    out.append({"n_gpus": n, "value": d["value"], "unit": d["unit"], "ms_per_step": d["ms_per_step"], "e2e": d["e2e"]["value"], "spp_per_s": d["spp_per_s"], "clocks": d["clocks"]})
for o in out: o["speedup_vs_1"] = o["value"]/out[0]["value"]; print(o["n_gpus"], round(o["value"],1), round(o["speedup_vs_1"],3))
json.dump(out, open("gpurun_out/scaling_1_2_4_8.json","w"), indent=1)
PY
