"""Phase timeline of ctrl_node_tc_kernel (CTA 0, warp 0) from a -DMACBF_TRACE build of the library:
   MACBF_B200_LIB=macbf-gnn_b200/libmacbf_b200_trace.so python macbf-gnn_b200/tools/node_trace.py"""
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
L = lib.load()
fn = L.macbf_debug_node_trace
fn.argtypes = [ctypes.c_void_p]
buf = np.zeros((1, 4096), dtype=np.uint64)
for _ in range(3):
    ops.controller_forward(s, g, idx, tf)     # the DUMP = false instantiation overwrites the table on every launch
torch.cuda.synchronize()
assert fn(buf.ctypes.data) == 0
names = {1: "entry", 2: "setup done", 3: "tile start", 4: "A0 stored", 5: "L1 done", 6: "E1 done+L2 issued", 7: "L2 done",
         8: "E2 done+L3 issued", 9: "L3 done", 10: "kernel end"}
ev = [(int(v) & 0x00ffffffffffffff, names[int(v) >> 56]) for v in buf[0] if int(v)]
t0 = ev[0][0]
prev = t0
for t, n in ev:
    print(f"{t - t0:9d} (+{t - prev:6d})  {n}")
    prev = t
