"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel + grid."""
import collections, csv, re, sys
lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
r = csv.reader(lines); hdr = next(r)
ki, vi, gi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
recs = [(row[ki], float(row[vi].replace(',', '')), row[gi]) for row in r]
agg = collections.defaultdict(lambda: [0, 0.0])
for k, v, g in recs:
    key = (re.sub(r'\(.*', '', k), g)
    agg[key][0] += 1; agg[key][1] += v
tot = sum(v for _, v, _ in recs)
print(f"{len(recs)} launches, total {tot/1e3:.1f} us (ncu per-launch times: cold-cache, serialised)")
for (n, g), (c, t) in sorted(agg.items(), key=lambda x: -x[1][1])[:30]:
    print(f"{t/1e3:9.1f} us {c:4d}x {t/c/1e3:8.1f} us/launch {100*t/tot:5.1f}%  {n[-40:]:40s} grid={g}")
