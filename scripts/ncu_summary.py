"""Summarise an .ncu-rep: headline metrics per kernel + warp-stall breakdown + hottest SASS lines.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [n_top]"""
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 14
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
want = ["Kernel Name", "gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__registers_per_thread",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum"]
for r in rows[2:]:
    print("-" * 100)
    for w in want:
        if w in hdr:
            i = hdr.index(w)
            print(f"{w:75s} {r[i][:60]} {rows[1][i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = re.split(r'(?m)^"Kernel Name",', src)
for blk in blocks[1:]:
    lines = list(csv.reader(io.StringIO('"Kernel Name",' + blk)))
    name = lines[0][1][:80]
    h = lines[1]
    isrc, isamp, iexe = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    stall = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
    body = [r for r in lines[2:] if len(r) > isamp and r[isamp].isdigit()]
    tot = sum(int(r[isamp]) for r in body) or 1
    agg = {}
    for r in body:
        for i in stall:
            if r[i].isdigit():
                agg[h[i][6:]] = agg.get(h[i][6:], 0) + int(r[i])
    print("=" * 100)
    print(name, "samples", tot)
    print("stalls:", {k: f"{100 * v / tot:.1f}%" for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v * 100 > tot})
    for r in sorted(body, key=lambda r: -int(r[isamp]))[:ntop]:
        st = {h[i][6:]: int(r[i]) for i in stall if r[i].isdigit() and int(r[i]) > 0}
        st = dict(sorted(st.items(), key=lambda x: -x[1])[:2])
        print(f"{100 * int(r[isamp]) / tot:5.1f}% exe={r[iexe]:>8s} {r[isrc].strip()[:64]:64s} {st}")
