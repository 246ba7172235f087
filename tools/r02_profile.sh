#!/bin/bash
# usage (GPU box): tools/r02_profile.sh <tag> [env...]  -- `ncu --set full` of one bounce>=1 closest launch and one shadow launch of a short bench command
TAG=$1; shift
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --spp 4 --no-cpu-baseline --no-roofline --no-stress"
env "$@" $CMD > gpurun_out/plain_$TAG.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain_$TAG.log; exit 1; }
env "$@" ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:k_traceILi.ELb0ELb0 -s 1 -c 1 -f -o gpurun_out/closest_full_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo ncu closest rc $?
