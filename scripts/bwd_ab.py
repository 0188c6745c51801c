"""A/B of the tensor-core CBF/controller backward against the FP32-FFMA kernels (selected per process by env)."""
import os, sys, subprocess, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path[:0] = [ROOT, os.path.join(ROOT, "macbf-gnn_b200")]
    import torch
    from macbf_b200 import ops
    from oracle import macbf_oracle as O
    B, N, k = 8, 1024, 12
    s, g = O.synthetic_swarm(B, N, 77)
    p_cbf, p_ctrl = O.init_params(5)
    cf = O.flatten_params(p_cbf, O.CBF_KEYS).cuda(); tf = O.flatten_params(p_ctrl, O.CTRL_KEYS).cuda()
    s, g = s.cuda(), g.cuda()
    idx = ops.neighbor_topk(s, k)
    gen = torch.Generator(device="cuda").manual_seed(1)
    gh = torch.randn(B, N, k, device="cuda", generator=gen) * (torch.rand(B, N, k, device="cuda", generator=gen) < 0.3)
    gu = torch.randn(B, N, 2, device="cuda", generator=gen)
    gs = torch.randn(B, N, 4, device="cuda", generator=gen)
    gp, gst = ops.cbf_backward(s, idx, cf, gh, want_grad_states=True)
    cp, cst, cg = ops.controller_backward(s, g, idx, tf, gu, gs, want_grad_states=True, want_grad_goals=True)
    torch.save({"gp": gp.cpu(), "gst": gst.cpu(), "cp": cp.cpu(), "cst": cst.cpu(), "cg": cg.cpu()}, sys.argv[2])
    sys.exit(0)
import torch
outs = {}
for name, env in (("tc", {}), ("ffma", {"MACBF_CBF_BWD_FFMA": "1", "MACBF_CTRL_BWD_FFMA": "1"})):
    f = f"/tmp/bwd_{name}.pt"
    subprocess.check_call([sys.executable, __file__, "child", f], env={**os.environ, **env})
    outs[name] = torch.load(f)
sys.path[:0] = [ROOT]
from oracle import macbf_oracle as O
def rel(a, b): return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))
off = 0
for key, shp in zip(O.CBF_KEYS, O.CBF_SHAPES):
    n = 1
    for d in shp: n *= d
    print(f"cbf {key:24s} {rel(outs['tc']['gp'][off:off+n], outs['ffma']['gp'][off:off+n]):.3e}")
    off += n
print("cbf grad_states", rel(outs['tc']['gst'], outs['ffma']['gst']))
off = 0
for key, shp in zip(O.CTRL_KEYS, O.CTRL_SHAPES):
    n = 1
    for d in shp: n *= d
    print(f"ctrl {key:32s} {rel(outs['tc']['cp'][off:off+n], outs['ffma']['cp'][off:off+n]):.3e}")
    off += n
print("ctrl grad_states", rel(outs['tc']['cst'], outs['ffma']['cst']), "grad_goals", rel(outs['tc']['cg'], outs['ffma']['cg']))
