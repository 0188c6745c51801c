"""CUDA-event timing of each stage of the forward step on the BASELINE configs[1] workload."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "macbf-gnn_b200")]
import torch
from macbf_b200 import ops
from oracle import macbf_oracle as O

B, N, k = int(os.environ.get("B", 64)), int(os.environ.get("N", 1024)), int(os.environ.get("K", 12))
s, g = O.synthetic_swarm(B, N, 3234)
p_cbf, p_ctrl = O.init_params(5)
cf = O.flatten_params(p_cbf, O.CBF_KEYS).cuda()
tf = O.flatten_params(p_ctrl, O.CTRL_KEYS).cuda()
s, g = s.cuda(), g.cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(name, fn, reps=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tot = 0.0
    for _ in range(reps):
        flush.zero_()
        e0.record(); fn(); e1.record(); e1.synchronize()
        tot += e0.elapsed_time(e1)
    print(f"{name:28s} {1e3 * tot / reps:9.1f} us")


idx = ops.neighbor_topk(s, k)
timeit("neighbor_topk", lambda: ops.neighbor_topk(s, k))
timeit("cbf_forward", lambda: ops.cbf_forward(s, idx, cf))
timeit("controller_forward", lambda: ops.controller_forward(s, g, idx, tf))
h = ops.cbf_forward(s, idx, cf)
u, sn, _ = ops.controller_forward(s, g, idx, tf)
timeit("loss_forward", lambda: ops.loss_forward(s, idx, h, h, u, g))
buf = ops.StepBuffers(B, N, 0, k, "cuda")
timeit("step_forward", lambda: ops.step_forward(s, g, cf, tf, buf))

if os.environ.get("BWD", "1") == "1":
    gh = torch.randn_like(h) * 1e-3
    gu = torch.randn_like(u) * 1e-3
    gs = torch.randn_like(sn) * 1e-3
    timeit("cbf_backward (params)", lambda: ops.cbf_backward(s, idx, cf, gh))
    timeit("cbf_backward (+d states)", lambda: ops.cbf_backward(s, idx, cf, gh, want_grad_states=True))
    timeit("controller_backward", lambda: ops.controller_backward(s, g, idx, tf, gu, gs))
    from macbf_b200.step import TrainStep
    ts = TrainStep(B, N, 0, k, "cuda")
    timeit("train forward_backward", lambda: ts.forward_backward(s, g, cf, tf))
