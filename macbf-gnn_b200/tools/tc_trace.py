"""Phase timeline of cbf_forward_tc_kernel (CTA 0): build the library with -DMACBF_TRACE into
build/libmacbf_b200_trace.so, run one launch on the configs[1] workload and print per-tile
phase offsets in SM cycles.   MACBF_B200_LIB=.../libmacbf_b200_trace.so python tools/tc_trace.py"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "macbf-gnn_b200")]
import numpy as np
import torch
from macbf_b200 import lib, ops
from oracle import macbf_oracle as O

s, g = O.synthetic_swarm(64, 1024, 3234)
p_cbf, _ = O.init_params(5)
cf = O.flatten_params(p_cbf, O.CBF_KEYS).cuda()
s = s.cuda()
idx = ops.neighbor_topk(s, 12)
for _ in range(3):
    ops.cbf_forward(s, idx, cf)
torch.cuda.synchronize()
L = lib.load()
buf = np.zeros((2, 4096), dtype=np.uint64)
fn = L.macbf_debug_tc_trace
fn.argtypes = [ctypes.c_void_p]
assert fn(buf.ctypes.data) == 0
names = {0: {1: "E2c0.start", 2: "E2c0.end", 3: "E2c1.start", 4: "E2c1.end", 5: "L1next.end", 6: "E3.start", 7: "E3.end", 8: "L1.features", 9: "L1.half0", 10: "L1.half1", 11: "L1.begin"},
         1: {1: "a2.ready", 2: "mma2.issued", 3: "a3[0].ready", 4: "a3[1].ready", 5: "mma3.issued"}}
c0, g0, c1, g1 = (int(x) for x in buf[1][4090:4094])
buf[1][4090:4094] = 0
ev = []
for who in (0, 1):
    for v in buf[who]:
        v = int(v)
        if v == 0:
            continue
        ev.append((v & 0x00ffffffffffffff, who, names[who][v >> 56]))
ev.sort()
t0 = ev[0][0]
last_a2 = None
for t, who, name in ev[:int(os.environ.get("NEV", 150))]:
    extra = ""
    if name == "a2.ready":
        if last_a2 is not None:
            extra = f"   <-- tile period {t - last_a2}"
        last_a2 = t
    print(f"{t - t0:9d}  {'worker0' if who == 0 else 'MMA    '}  {name}{extra}")

first_a2 = min(t for t, who, n in ev if n == "a2.ready")
print(f"kernel (CTA 0): {c1 - c0} SM cycles in {g1 - g0} ns -> SM clock {(c1 - c0) / max(g1 - g0, 1):.3f} GHz; "
      f"entry -> first a2.ready {first_a2 - c0} cycles; tiles {sum(1 for t, w, n in ev if n == 'a2.ready')}")
