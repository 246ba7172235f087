#!/bin/bash
# usage (GPU box): tools/profile_src.sh <tag> [kernel-regex] [skip]  -- ncu source counters for one launch of a trace kernel
TAG=${1:-x}; K=${2:-k_trace_closest}; S=${3:-5}
CMD="python bench.py --steps 1 --warmup 1 --spp 4 --no-cpu-baseline --no-roofline --no-stress"
$CMD > gpurun_out/plain_src_$TAG.log 2>&1 && ncu --section SourceCounters --section InstructionStats --section WarpStateStats --clock-control none --import-source on -k regex:$K -s $S -c 1 -o gpurun_out/prof_src_$TAG $CMD > gpurun_out/ncu_src_$TAG.log 2>&1
echo rc $?
