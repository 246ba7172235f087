#!/bin/bash
# usage (GPU box): tools/r02_evidence.sh  -- everything profiles/r02_* is made from, each ncu pass only after its command ran clean
mkdir -p gpurun_out/ev
E=gpurun_out/ev
CMD="python bench.py --steps 1 --warmup 1 --spp 4 --samples-per-launch 4 --no-cpu-baseline --no-stress --no-strong"
$CMD > $E/plain_atrium.json 2> $E/plain_atrium.err || { echo plain failed; tail -5 $E/plain_atrium.err; exit 1; }
# 1. launch list of the whole command
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $E/launches.csv $CMD > $E/ncu_launches.log 2>&1; echo launches rc $?
# 2. light counters over the 4 closest + 4 shadow launches of one step (warm-up step: identical rays every step)
M=gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active
ncu --metrics $M --clock-control none --kernel-name-base mangled -k regex:k_traceILi.ELb -c 8 --csv --log-file $E/trace_light.csv $CMD > $E/ncu_light.log 2>&1; echo light rc $?
# 3. full captures: closest bounce 1, shadow bounce 1, shade bounce 1
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:k_traceILi.ELb0ELb0 -s 1 -c 1 -f -o $E/atrium_closest_full $CMD > $E/ncu_c.log 2>&1; echo closest rc $?
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:k_traceILi.ELb1ELb0 -s 1 -c 1 -f -o $E/atrium_shadow_full $CMD > $E/ncu_s.log 2>&1; echo shadow rc $?
ncu --set full --clock-control none --import-source on -k regex:k_shade -s 1 -c 1 -f -o $E/atrium_shade_full $CMD > $E/ncu_h.log 2>&1; echo shade rc $?
# 4. district (configs[3]): bench line with its roofline block, then one full capture of a bounce-1 closest launch
DC="python bench.py --workload district --steps 2 --warmup 2"
timeout 900 $DC > $E/bench_district.json 2> $E/bench_district.err; echo district rc $?
ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:k_traceILi.ELb0ELb0 -s 1 -c 1 -f -o $E/district_closest_full python bench.py --workload district --steps 1 --warmup 0 > $E/ncu_d.log 2>&1; echo district ncu rc $?
