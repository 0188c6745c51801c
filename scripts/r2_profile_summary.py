"""Summaries for profiles/: per-kernel table of an ncu launch list with metrics, and headline metrics of a --set full report.
usage: python scripts/r2_profile_summary.py launches <csv>  |  full <ncu-rep>  |  json <out.json> <ncu-rep> [<ncu-rep> ...]
(json: per-kernel DRAM bytes / tensor-pipe activity under bench.py's kernel labels, read by bench.py for `roofline.traffic`)"""
import collections
import csv
import io
import json
import subprocess
import sys

mode, path = sys.argv[1], sys.argv[2]


def bench_label(name):
    """ncu kernel name -> the label bench.py's live table uses (MACBF_CHECK_LAUNCH strings)."""
    import re
    name = name.replace("void ", "").replace("macbf::", "").strip()
    m = re.match(r"(\w+)(<.*>)?", name)
    base, targs = m.group(1), (m.group(2) or "")
    first = re.sub(r"[<>()a-z ]", "", targs).split(",")[0] if targs else ""
    if base == "cbf_forward_tc_kernel":
        return base + {"0": "", "1": "<dump>", "2": "<record>"}.get(first, targs)
    if base == "ctrl_node_tc_kernel":
        return base + {"0": "", "1": "<dump>"}.get(first, targs)
    if base == "wgrad_tc_kernel":
        return base + {"1": "<ctrl>", "2": "<cbf>", "3": "<node>"}.get(first, targs)
    return base


if mode == "json":
    out = {"workload": [1024, 64, 12, 0],
           "how": "ncu --set full --clock-control none, one launch per kernel of `python bench.py` (scripts/r2_profiles.sh); per launch",
           "kernels": {}}
    for rep in sys.argv[3:]:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr = rows[0]
        col = {w: hdr.index(w) for w in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
                                          "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed")}
        unit = rows[1]

        def val(r, w):
            v = float(r[col[w]].replace(",", ""))
            u = unit[col[w]].lower()
            return v * {"gbyte": 1e9, "mbyte": 1e6, "kbyte": 1e3, "byte": 1.0, "ms": 1e3, "us": 1.0, "ns": 1e-3, "%": 1.0, "msecond": 1e3,
                        "usecond": 1.0, "nsecond": 1e-3}.get(u, 1.0)
        for r in rows[2:]:
            lab = bench_label(r[col["Kernel Name"]].split("(")[0] if "<" not in r[col["Kernel Name"]] else r[col["Kernel Name"]])
            if lab in out["kernels"]:
                continue
            rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
            out["kernels"][lab] = {"dram_bytes": rd + wr, "dram_read": rd, "dram_write": wr,
                                   "tensor_active_pct": val(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"),
                                   "duration_us_under_ncu": val(r, "gpu__time_duration.sum")}
    json.dump(out, open(path, "w"), indent=1)
    print(json.dumps({k: round(v["duration_us_under_ncu"], 1) for k, v in out["kernels"].items()}))
elif mode == "launches":
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
