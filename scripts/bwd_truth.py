"""Controller / CBF backward (tensor-core and FFMA paths) against an fp64 autograd evaluation of the oracle:
separates kernel error from the inherent ReLU-kink sensitivity of fp32 gradients."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "macbf-gnn_b200")]
import torch
from macbf_b200 import ops
from oracle import macbf_oracle as O

B, N, k = 2, 1024, 12
s, g = O.synthetic_swarm(B, N, 77)
p_cbf, p_ctrl = O.init_params(5)
gen = torch.Generator().manual_seed(1)
gu = torch.randn(B, N, 2, generator=gen)
gs = torch.randn(B, N, 4, generator=gen)
gh = torch.randn(B, N, k, generator=gen) * (torch.rand(B, N, k, generator=gen) < 0.3)


def truth(dtype):
    pc = {n: t.to(dtype).clone().requires_grad_(True) for n, t in p_cbf.items()}
    pt = {n: t.to(dtype).clone().requires_grad_(True) for n, t in p_ctrl.items()}
    tot = 0
    for b in range(B):
        sb, gb = s[b].to(dtype), g[b].to(dtype)
        u = O.controller_forward(pt, sb, gb)
        sn = O.euler_step(sb, u)
        h = O.cbf_forward(pc, sb)
        tot = tot + (u * gu[b].to(dtype)).sum() + (sn * gs[b].to(dtype)).sum() + (h[:, 0, :] * gh[b].to(dtype)).sum()
    tot.backward()
    return (torch.cat([pc[n].grad.reshape(-1) for n in O.CBF_KEYS]).float(),
            torch.cat([pt[n].grad.reshape(-1) for n in O.CTRL_KEYS]).float())


t64 = truth(torch.float64)
t32 = truth(torch.float32)


def gpu():
    cf = O.flatten_params(p_cbf, O.CBF_KEYS).cuda(); tf = O.flatten_params(p_ctrl, O.CTRL_KEYS).cuda()
    sd, gd = s.cuda(), g.cuda()
    idx = ops.neighbor_topk(sd, k)
    gp, _ = ops.cbf_backward(sd, idx, cf, gh.cuda())
    cp, _, _ = ops.controller_backward(sd, gd, idx, tf, gu.cuda(), gs.cuda())
    return gp.cpu(), cp.cpu()


def rel(a, b): return float((a - b).abs().max() / b.abs().max())


tc = gpu()
os.environ["MACBF_CBF_BWD_FFMA"] = "1"; os.environ["MACBF_CTRL_BWD_FFMA"] = "1"
ff = gpu()
for name, i in (("cbf", 0), ("ctrl", 1)):
    print(f"{name}: torch fp32 vs fp64 {rel(t32[i], t64[i]):.2e} | tensor-core vs fp64 {rel(tc[i], t64[i]):.2e} | "
          f"FFMA vs fp64 {rel(ff[i], t64[i]):.2e} | tensor-core vs FFMA {rel(tc[i], ff[i]):.2e}")
