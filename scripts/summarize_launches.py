"""Summarise an ncu --metrics gpu__time_duration.sum --csv launch list by kernel."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
start = rows.index(hdr)
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[start + 1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    n = r[ki].split("(")[0][:64]
    agg[n][0] += 1
    agg[n][1] += v
tot = sum(v[1] for v in agg.values())
print(f"{'kernel':66s} {'n':>4s} {'total_us':>10s} {'avg_us':>9s} {'share':>6s}")
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{n:66s} {c:4d} {t / 1e3:10.1f} {t / c / 1e3:9.1f} {100 * t / tot:5.1f}%")
