#!/bin/bash
# usage (on the GPU box): tools/quick_gpu.sh <tag>   -- GPU tests, a 2-step bench, and a light ncu pass over the trace kernels
TAG=${1:-x}
mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5)
timeout 900 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo bench exit $?
python - <<PY
import json
d=json.load(open("gpurun_out/bench_$TAG.json"))
print(d["tlas"]); print("Mrays/s", d["value"], "ms/step", d["ms_per_step"], d["breakdown"]["stage_ms_last_step"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"])
PY
tail -3 gpurun_out/bench_$TAG.err
CMD="python bench.py --steps 1 --warmup 1 --spp 4 --no-cpu-baseline --no-roofline --no-stress"
$CMD > gpurun_out/plain_$TAG.log 2>&1 && ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:k_trace -s 8 -c 8 --csv --log-file gpurun_out/light_$TAG.csv $CMD > gpurun_out/ncu_$TAG.log 2>&1
echo ncu rc $?
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/light_$TAG.csv")) if len(r)>8]
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ii=h.index("ID")
cur={}
for r in rows[1:]:
    cur.setdefault((r[ii], r[ki].split("(")[0][-20:]), {})[r[mi].split(".")[0][-40:]]=r[vi]
for k,v in cur.items(): print(k, v)
PY
