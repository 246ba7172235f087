"""Aggregates tools/hotspots.sh output lines (stdin, `... <top>` large) into named source ranges. usage: hotspots.sh rep kernel 100000 | hot_ranges.py file:lo-hi=name ..."""
import sys, re
ranges = []
for a in sys.argv[1:]:
    loc, name = a.split("=")
    f, r = loc.split(":")
    lo, hi = map(int, r.split("-"))
    ranges.append((f, lo, hi, name))
agg = {}
tot_w = tot_s = 0.0
for line in sys.stdin:
    m = re.match(r"(\S+)\s*:\s*(\d+)\s+sass\s+(\d+)\s+warp-inst\s+([\d.]+)%\s+lanes\s+([\d.]+)\s+samples\s+(\d+)", line)
    if not m:
        continue
    f, ln, sass, w, lanes, smp = m.group(1), int(m.group(2)), int(m.group(3)), float(m.group(4)), float(m.group(5)), int(m.group(6))
    name = next((n for (rf, lo, hi, n) in ranges if rf == f and lo <= ln <= hi), f"other:{f}")
    a = agg.setdefault(name, [0.0, 0.0, 0, 0])
    a[0] += w; a[1] += w * lanes; a[2] += smp; a[3] += sass
    tot_w += w; tot_s += smp
for n, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{n:34s} warp-inst {a[0]:5.1f}%  lanes {a[1]/max(a[0],1e-9):5.1f}  samples {a[2]/max(tot_s,1)*100:5.1f}%  sass {a[3]}")
