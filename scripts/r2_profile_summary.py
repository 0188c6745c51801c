"""Summaries for profiles/: per-kernel table of an ncu launch list with metrics, and headline metrics of a --set full report.
usage: python scripts/r2_profile_summary.py launches <csv>  |  full <ncu-rep>"""
import collections
import csv
import io
import json
import subprocess
import sys

mode, path = sys.argv[1], sys.argv[2]
if mode == "launches":
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = next(r for r in rows if "Kernel Name" in r)
    st = rows.index(hdr)
    ki, mi, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
    agg = collections.defaultdict(lambda: collections.defaultdict(list))
    for r in rows[st + 1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        agg[r[ki].split("(")[0][:60]][r[mi]].append(v)
    tot = sum(sum(m.get("gpu__time_duration.sum", [0])) for m in agg.values())
    print(f"{'kernel':62s} {'n':>4s} {'avg_us':>8s} {'share':>6s} {'rd_MB':>8s} {'wr_MB':>8s} {'tensor%':>7s}")
    for n, m in sorted(agg.items(), key=lambda x: -sum(x[1].get("gpu__time_duration.sum", [0]))):
        d = m.get("gpu__time_duration.sum", [0])
        g = lambda k: sum(m.get(k, [0])) / len(d)
        print(f"{n:62s} {len(d):4d} {sum(d) / len(d) / 1e3:8.1f} {100 * sum(d) / tot:5.1f}% {g('dram__bytes_read.sum') / 1e6:8.2f} "
              f"{g('dram__bytes_write.sum') / 1e6:8.2f} {g('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed'):7.1f}")
else:
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = rows[0]
    want = ["gpu__time_duration.sum", "launch__registers_per_thread", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_bytes.sum"]
    seen = set()
    out = {}
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0]
        if name in seen:
            continue
        seen.add(name)
        print("-" * 100)
        print(name)
        d = {}
        for w in want:
            if w in hdr:
                i = hdr.index(w)
                print(f"  {w:72s} {r[i]:>16s} {rows[1][i]}")
                d[w] = (r[i], rows[1][i])
        out[name] = d
