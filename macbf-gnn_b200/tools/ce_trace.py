"""Phase timeline of ctrl_edge_tc2_kernel (CTA 0) from a -DMACBF_TRACE build of the library:
   MACBF_B200_LIB=macbf-gnn_b200/libmacbf_b200_trace.so python macbf-gnn_b200/tools/ce_trace.py"""
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
if hasattr(L, "macbf_debug_ce2_trace"):
    buf = np.zeros((4, 4096), dtype=np.uint64)
    fn = L.macbf_debug_ce2_trace
    fn.argtypes = [ctypes.c_void_p]
    assert fn(buf.ctypes.data) == 0
    names = {0: {1: "row  iter.start", 2: "row  d1.ready", 3: "row  E1.stored", 4: "row  a2.arrived", 5: "row  feat.done", 10: "KERNEL ENTRY", 11: "SETUP DONE", 12: "KERNEL END", 6: "row    F.arrived", 7: "row    sj.issued", 8: "row    F.geom", 9: "row    F.stored"},
             1: {1: "                    MMA0 a2.ready", 2: "                    MMA0 L2.issued", 3: "                    MMA0 L1.issued"},
             3: {1: "                    MMA1 a2.ready", 2: "                    MMA1 L2.issued", 3: "                    MMA1 L1.issued"},
             2: {1: "                                        pool iter.start", 2: "                                        pool d2.ready",
                 3: "                                        pool loaded"}}
    ev = []
    for who in range(4):
        for v in buf[who]:
            v = int(v)
            if v:
                ev.append((v & 0x00ffffffffffffff, names[who][v >> 56]))
    ev.sort()
    skip = int(os.environ.get("SKIP", 150))
    t0 = ev[skip][0]
    for t, name in ev[skip:skip + int(os.environ.get("NEV", 160))]:
        print(f"{t - t0:9d}  {name}")
    print("first events:")
    for t, name in ev[:40]:
        print(f"{t - ev[0][0]:9d}  {name}")
    print("last events:")
    for t, name in ev[-12:]:
        print(f"{t - ev[0][0]:9d}  {name}")
    a2 = [t for t, n in ev if "a2.ready" in n]
    print("tiles", len(a2), "mean period", (a2[-1] - a2[0]) / max(len(a2) - 1, 1))

# per-kernel event times of controller_forward, warm L2 vs flushed L2
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for label, do_flush in (("warm", False), ("flushed", True)):
    acc = {}
    for _ in range(10):
        if do_flush:
            flush.zero_()
        torch.cuda.synchronize()
        ops.timing_begin()
        ops.controller_forward(s, g, idx, tf)
        for name, ms in ops.timing_end():
            acc.setdefault(name, []).append(ms)
    print(label, {k: round(1e3 * sum(v) / len(v), 1) for k, v in acc.items()}, "us")
