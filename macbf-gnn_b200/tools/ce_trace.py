"""Phase timeline of ctrl_edge_tc_kernel (CTA 0), MACBF_TRACE build (see tools/tc_trace.py)."""
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
_, p_ctrl = O.init_params(5)
tf = O.flatten_params(p_ctrl, O.CTRL_KEYS).cuda()
s, g = s.cuda(), g.cuda()
idx = ops.neighbor_topk(s, 12)
for _ in range(3):
    ops.controller_forward(s, g, idx, tf)
torch.cuda.synchronize()
L = lib.load()
buf = np.zeros((3, 4096), dtype=np.uint64)
fn = L.macbf_debug_ce_trace
fn.argtypes = [ctypes.c_void_p]
assert fn(buf.ctypes.data) == 0
wn = {1: "d2.ready", 2: "E2h0.end", 3: "bar.out", 4: "(h0)", 5: "pool0.end", 6: "E2h1.end", 7: "bar.out", 8: "E1next.end", 9: "pool1.end", 10: "feat.end"}
names = {0: {k: "w0 " + v for k, v in wn.items()}, 2: {k: "        w4 " + v for k, v in wn.items()},
         1: {1: "MMA a2.ready", 2: "MMA mma2.issued", 3: "MMA mma1.issued"}}
ev = []
for who in (0, 1, 2):
    for v in buf[who]:
        v = int(v)
        if v:
            ev.append((v & 0x00ffffffffffffff, names[who][v >> 56]))
ev.sort()
t0 = ev[0][0]
last = None
for t, name in ev[:int(os.environ.get("NEV", 200))]:
    extra = ""
    if name == "MMA a2.ready":
        if last is not None:
            extra = f"   <-- tile period {t - last}"
        last = t
    print(f"{t - t0:9d}  {name}{extra}")
