"""Rollout timing: fused macbf_rollout_forward vs the per-step API with the reference's host-side
stop test (one .item() sync per step, train.py:78).  Usage: python scripts/rollout_time.py [B N T]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "macbf-gnn_b200"))
import torch
from macbf_b200 import ops
import oracle.macbf_oracle as O

B, N, T = (int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (64, 1024, 16)
_, p = O.init_params(0)
flat = O.flatten_params(p, O.CTRL_KEYS).cuda()
s, g = O.synthetic_swarm(B, N, 1)
s, g = s.cuda(), g.cuda()


def per_step():
    cur = s
    for _ in range(T):
        idx = ops.neighbor_topk(cur, 12)
        u, cur, _ = ops.controller_forward(cur, g, idx, flat)
        if bool(torch.mean(torch.norm(cur[..., :2] - g, dim=-1)) < 0.07):
            break
    return cur


def fused():
    return ops.rollout_forward(s, g, flat, T)


for name, fn in (("per_step_host_stop", per_step), ("fused_device_stop", fused)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 10
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    print(f"{name}: {dt * 1e3:.3f} ms per rollout of {T} steps, {dt / T * 1e6:.1f} us/step, "
          f"{B * N * T / dt / 1e6:.1f} M agent-steps/s")
