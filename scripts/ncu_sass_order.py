"""Print the SASS of one kernel of an .ncu-rep in program order with stall samples (lines above a threshold or of
interesting opcodes).  usage: python scripts/ncu_sass_order.py rep [min_samples] [kernel-index]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]; thr = int(sys.argv[2]) if len(sys.argv) > 2 else 10; ki = int(sys.argv[3]) if len(sys.argv) > 3 else 0
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
blocks = re.split(r'(?m)^"Kernel Name",', src)[1:]
lines = list(csv.reader(io.StringIO('"Kernel Name",' + blocks[ki])))
h = lines[1]
isrc, isamp, iexe = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
stall = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
body = [r for r in lines[2:] if len(r) > isamp and r[isamp].isdigit()]
tot = sum(int(r[isamp]) for r in body) or 1
print(lines[0][1][:100], "samples", tot)
key = re.compile(r"LDTM|STTM|UTCHMMA|SYNCS|UTCBAR|BAR\.|LDG|STG|STS|LDS|FENCE|WARPSYNC|ELECT|EXIT")
acc = 0
for n, r in enumerate(body):
    s = int(r[isamp]); acc += s
    if s >= thr or key.search(r[isrc]):
        st = {h[i][6:]: int(r[i]) for i in stall if r[i].isdigit() and int(r[i]) > 0}
        st = dict(sorted(st.items(), key=lambda x: -x[1])[:2])
        print(f"{n:5d} {100*s/tot:5.1f}% cum {100*acc/tot:5.1f}% exe={r[iexe]:>7s} {r[isrc].strip()[:70]:70s} {st}")
