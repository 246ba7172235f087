"""Joins `ncu --page source --csv` (per-SASS-instruction counters) with `nvdisasm -g -c` line info of the same
kernel and prints the hottest source lines.  usage: sass_hotspots.py <src_sass.csv> <kernel.sass> [top]"""
import csv, re, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ie, te = h.index("Instructions Executed"), h.index("Thread Instructions Executed")
si = h.index("Source")
samp = h.index("# Samples") if "# Samples" in h else None
body = []
for r in rows[hi + 1:]:  # the first kernel table only
    if not r or r[0] in ("Kernel Name", "Address"):
        break
    body.append(r)
inst = [(r[si].strip(), int(r[ie]), int(r[te]), int(r[samp]) if samp is not None else 0) for r in body if len(r) > te]
loc, cur = [], ("?", 0)
for line in open(sys.argv[2]):
    m = re.search(r'//## File "(.*)", line (\d+)', line)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.search(r"/\*[0-9a-f]{4,}\*/", line):
        loc.append((cur, line.split("*/", 1)[1].strip()))
print("csv instr", len(inst), "sass instr", len(loc))
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
for (src, ex, tex, sm), (l, txt) in zip(inst, loc):
    a = agg[l]; a[0] += ex; a[1] += tex; a[2] += sm; a[3] += 1
tot = sum(a[0] for a in agg.values()); tt = sum(a[1] for a in agg.values())
print(f"total warp-inst {tot/1e9:.2f} G, thread-inst {tt/1e9:.2f} G, avg lanes {tt/max(tot,1):.2f}")
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
key = 2 if len(sys.argv) > 4 and sys.argv[4] == "samples" else 0
print("total samples", sum(a[2] for a in agg.values()))
for l, a in sorted(agg.items(), key=lambda kv: -kv[1][key])[:top]:
    print(f"{l[0]:28s}:{l[1]:4d}  sass {a[3]:4d}  warp-inst {a[0]/tot*100:5.1f}%  lanes {a[1]/max(a[0],1):5.1f}  samples {a[2]}")
