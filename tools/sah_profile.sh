#!/bin/bash
# usage (GPU box): tools/sah_profile.sh -- ncu launch list (time + DRAM bytes) of the BLAS builder kernels, after a plain run
mkdir -p gpurun_out
CMD="python tools/blas_build_probe.py"
PYTHONPATH=. $CMD > gpurun_out/blas_probe.log 2>&1 || { tail -5 gpurun_out/blas_probe.log; exit 1; }
tail -1 gpurun_out/blas_probe.log
PYTHONPATH=. ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'k_sah|k_collapse|k_blas_gather|k_pad_and_keys|k_emit_tris' -c 3000 --csv --log-file gpurun_out/blas_build_launches.csv $CMD > gpurun_out/ncu_blas.log 2>&1
echo ncu rc $?
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/blas_build_launches.csv")) if len(r)>8]
h=rows[0]; ki=h.index("Kernel Name"); mi=h.index("Metric Name"); vi=h.index("Metric Value"); ui=h.index("Metric Unit")
t=collections.defaultdict(lambda: [0,0.0,0.0])
for r in rows[1:]:
    k=r[ki].split("(")[0]; v=float(r[vi].replace(",",""))
    if r[mi]=="gpu__time_duration.sum":
        t[k][0]+=1; t[k][1]+= v/1e3 if r[ui] in ("ns","nsecond") else (v if r[ui] in ("us","usecond") else v*1e3)
    else:
        t[k][2]+= v * {"byte":1,"Kbyte":1e3,"Mbyte":1e6,"Gbyte":1e9}.get(r[ui],1)
print("kernel,launches,total_us,dram_MB")
for k,(n,us,b) in sorted(t.items(), key=lambda x:-x[1][1]): print(f"{k},{n},{us:.1f},{b/1e6:.1f}")
PY
