"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel and grid."""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H, data = rows[hdr], rows[hdr + 1:]
ki, vi, ui, gi = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit"), H.index("Grid Size")
agg, tot = collections.OrderedDict(), 0.0
for r in data:
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("<unnamed>::", "")
    v = float(r[vi].replace(",", ""))
    v = v / 1000 if r[ui] == "ns" else v * 1000 if r[ui] == "ms" else v
    a = agg.setdefault((name, r[gi]), [0, 0.0])
    a[0] += 1; a[1] += v; tot += v
print(f"total {tot:.1f} us over {len(data)} launches (cold-cache, serialised: compare shares)")
for (name, grid), (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{t:9.1f} us {n:4d}x avg {t / n:7.2f} us {100 * t / tot:5.1f}%  {name[:64]} grid={grid}")
