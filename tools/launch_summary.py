"""usage: python tools/launch_summary.py <ncu launch-list csv> > summary.csv  -- per-kernel launches, total time, share"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 8]
h = rows[0]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    name = r[ki].split("(")[0].replace("prt::", "")
    tot[name] += float(r[vi].replace(",", "")) / 1e6
    cnt[name] += 1
total = sum(tot.values())
print("kernel,launches,total_ms,share_pct")
for k, v in tot.most_common():
    print(f"{k},{cnt[k]},{v:.3f},{100 * v / total:.1f}")
