"""Turns gpurun_out/ev/* (tools/r02_evidence.sh, run on the B200) into the committed summaries under profiles/r02_*.
usage (here): python tools/r02_make_profiles.py"""
import csv, json, os, subprocess, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
E, P = os.path.join(ROOT, "gpurun_out", "ev"), os.path.join(ROOT, "profiles")

def ncu_csv(rep, page):
    return subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout

# launch list + per-kernel summary
rows = [r for r in csv.reader(open(os.path.join(E, "launches.csv"))) if len(r) > 8]
open(os.path.join(P, "r02_launches.csv"), "w").write(open(os.path.join(E, "launches.csv")).read())
h = rows[0]; ki, vi = h.index("Kernel Name"), h.index("Metric Value")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    n = r[ki].split("(")[0].replace("prt::", "")
    tot[n] += float(r[vi].replace(",", "")) / 1e6; cnt[n] += 1
total = sum(tot.values())
with open(os.path.join(P, "r02_launches_summary.csv"), "w") as f:
    f.write("kernel,launches,total_ms,share_pct\n")
    for k, v in tot.most_common():
        f.write(f"{k},{cnt[k]},{v:.3f},{100 * v / total:.1f}\n")

# light counters over the closest / shadow launches of one step -> warp instructions per ray
rows = [r for r in csv.reader(open(os.path.join(E, "trace_light.csv"))) if len(r) > 8]
open(os.path.join(P, "r02_trace_light.csv"), "w").write(open(os.path.join(E, "trace_light.csv")).read())
h = rows[0]; ki, mi, vi, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
per = collections.OrderedDict()
for r in rows[1:]:
    per.setdefault(r[ii], {"kernel": r[ki]})[r[mi]] = float(r[vi].replace(",", ""))
plain = json.load(open(os.path.join(E, "plain_atrium.json")))
cc = plain["breakdown"]["counters_per_step"]
out = {"source": "tools/r02_evidence.sh: ncu --metrics ... over the 4 closest + 4 shadow launches of one 4-spp step (1080p atrium, 4 bounces)",
       "counters_per_step": {k: cc[k] for k in ("closestRays", "shadowRays")}}
for name, tag, rays in (("closest", "k_trace<2, 0,", cc["closestRays"]), ("shadow", "k_trace<2, 1,", cc["shadowRays"])):
    sel = [v for v in per.values() if tag in v["kernel"]]
    inst = sum(v["smsp__inst_executed.sum"] for v in sel)
    w = lambda m: sum(v[m] * v["gpu__time_duration.sum"] for v in sel) / sum(v["gpu__time_duration.sum"] for v in sel)
    out[name] = {"launches": len(sel), "warp_inst": inst, "warp_inst_per_ray": inst / rays,
                 "lanes": sum(v["smsp__thread_inst_executed_per_inst_executed.ratio"] * v["smsp__inst_executed.sum"] for v in sel) / inst,
                 "issue_active_pct": w("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                 "l1_data_pipe_pct": w("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
                 "l1_hit_pct": w("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": w("lts__t_sector_hit_rate.pct"),
                 "occupancy_pct": w("sm__warps_active.avg.pct_of_peak_sustained_active"),
                 "dram_bytes_per_launch": sum(v["dram__bytes_read.sum"] + v["dram__bytes_write.sum"] for v in sel) / len(sel),
                 "ms": sum(v["gpu__time_duration.sum"] for v in sel) / 1e6}
json.dump(out, open(os.path.join(P, "r02_trace_kernel_counters.json"), "w"), indent=1)
print(json.dumps(out, indent=1))

# full captures
for rep, name in (("atrium_closest_full", "r02_atrium_k_trace_closest_full"), ("atrium_shadow_full", "r02_atrium_k_trace_shadow_full"),
                  ("atrium_shade_full", "r02_atrium_k_shade_full"), ("district_closest_full", "r02_district_k_trace_closest_full")):
    path = os.path.join(E, rep + ".ncu-rep")
    if not os.path.exists(path):
        print("missing", path); continue
    open(os.path.join(P, name + "_details.csv"), "w").write(ncu_csv(path, "details"))
    open(os.path.join(P, name + "_raw.csv"), "w").write(ncu_csv(path, "raw"))
for src, dst in (("bench_district.json", "r02_bench_district_4k.json"),):
    open(os.path.join(P, dst), "w").write(open(os.path.join(E, src)).read())
