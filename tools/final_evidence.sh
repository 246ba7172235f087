#!/bin/bash
# usage (GPU box): tools/final_evidence.sh 1|2
#   1: GPU tests, the default bench (-> gpurun_out/bench_final.json), the reference arm, then the ncu launch list of a
#      short bench command (after that command exited 0 without ncu)
#   2: one `ncu --set full` capture of a bounce>=1 k_trace_closest launch of the same short command
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --spp 4 --no-cpu-baseline --no-roofline --no-stress"
if [ "$1" = "1" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
  timeout 900 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo bench rc $?
  timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference_final.json 2>> gpurun_out/bench_final.err; echo reference rc $?
  $CMD > gpurun_out/plain_final.log 2>&1 || exit 1
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_final.csv $CMD > gpurun_out/ncu_launches.log 2>&1
  echo ncu rc $?
else
  $CMD > gpurun_out/plain_full.log 2>&1 || exit 1
  ncu --set full --clock-control none --import-source on -k regex:k_trace_closest -s 5 -c 1 -f -o gpurun_out/closest_full_final $CMD > gpurun_out/ncu_full.log 2>&1
  echo ncu rc $?
fi
