"""Probe: the fused forward step replayed from a CUDA graph vs launched directly (configs[1])."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "macbf-gnn_b200")]
import torch
from macbf_b200 import ops
from oracle import macbf_oracle as O

B, N, K = 64, 1024, 12
s, g = O.synthetic_swarm(B, N, 1234)
p_cbf, p_ctrl = O.init_params(0)
cf, tf = O.flatten_params(p_cbf, O.CBF_KEYS).cuda(), O.flatten_params(p_ctrl, O.CTRL_KEYS).cuda()
s, g = s.cuda(), g.cuda()
buf = ops.StepBuffers(B, N, 0, K, "cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

def step():
    ops.step_forward(s, g, cf, tf, buf)

def timeit(fn, n=30):
    for _ in range(5):
        fn()
    ts = []
    for _ in range(n):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return sum(ts) / len(ts), ts[len(ts) // 2]

print("direct  mean/median ms", timeit(step))
ref = buf.scalars.clone()
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        step()
torch.cuda.synchronize()
print("graph   mean/median ms", timeit(gr.replay))
print("scalars equal:", torch.equal(ref, buf.scalars))
