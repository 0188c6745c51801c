"""GPU parity (backward / training step): parameter gradients of the REPAIRED step against the
oracle's torch-autograd gradients (golden fixtures), state gradients against autograd on the CPU
oracle, and the drop-in modules' autograd against the same goldens.

Tolerance: gradients are sums over up to ~12k edges of fp32 products; 5e-5 scale-relative
(relative to the largest gradient entry of the tensor) is used for them, 1e-5 for losses.
"""
import numpy as np
import pytest
import torch

import os

from helpers import (GOLDEN, RTOL, assert_close, dev, load_npz, oracle_cases, oracle_params, oracle_step_on_device_graph,
                     rel_err)

pytestmark = pytest.mark.gpu
GTOL = 5e-5


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from macbf_b200 import ops as _ops
    return _ops


def _inputs(z):
    s = dev(z["states"]).unsqueeze(0)
    g = dev(z["goals"]).unsqueeze(0)
    obs = dev(z["obstacles"]).unsqueeze(0) if z["obstacles"].shape[0] else None
    return s, g, obs


@pytest.mark.parametrize("case", oracle_cases())
def test_fused_train_step_gradients(ops, case):
    from macbf_b200 import step
    z = load_npz(case)
    s, g, obs = _inputs(z)
    k = int(z["k"])
    N = s.shape[1]
    M = 0 if obs is None else obs.shape[1]
    ts = step.TrainStep(1, N, M, k, "cuda")
    scal, grad = ts.forward_backward(s, g, dev(z["cbf_flat"]), dev(z["ctrl_flat"]), obs)
    torch.cuda.synchronize()
    if (ts.buf.idx_next[0].cpu().numpy() == z["idx_next"]).all():
        ref = {n: torch.from_numpy(z[n]) for n in ("total", "grad_cbf", "grad_ctrl")}
    else:   # a near-tie row of the S' graph flipped (u differs by 1 ulp): compare on the device's graph, never skip
        p_cbf, p_ctrl = oracle_params(z)
        ref, _ = oracle_step_on_device_graph(p_cbf, p_ctrl, torch.from_numpy(z["states"]), torch.from_numpy(z["goals"]),
                                             torch.from_numpy(z["obstacles"]) if z["obstacles"].shape[0] else None,
                                             k, ts.buf.idx_next[0].cpu().numpy())
    assert abs(float(scal[0, 9]) - float(ref["total"])) <= 2 * RTOL * abs(float(ref["total"]))
    gc = grad[:ops.CBF_PARAMS].cpu().numpy()
    gt = grad[ops.CBF_PARAMS:].cpu().numpy()
    assert_close(gc, ref["grad_cbf"].numpy(), GTOL, "grad_cbf")
    assert_close(gt, ref["grad_ctrl"].numpy(), GTOL, "grad_ctrl")
    # the same step pinned to the mechanically repaired reference (oracle/repairs.py), where it has semantics
    rr = os.path.join(GOLDEN, case.replace("oracle_", "ref_repaired_"))
    if os.path.isfile(rr) and (ts.buf.idx_next[0].cpu().numpy() == z["idx_next"]).all():
        r = load_npz(os.path.basename(rr))
        assert np.array_equal(ts.buf.ttc_mask[0].cpu().numpy().astype(bool).reshape(-1), r["ttc_mask"].reshape(-1))
        sc = scal[0].cpu().numpy()
        for q in range(9):
            assert abs(float(sc[q]) - float(r["scalars"][q])) <= 2 * RTOL * max(abs(float(r["scalars"][q])), 1e-3), q
        assert_close(gc, r["grad_cbf"], GTOL, "grad_cbf vs repaired reference")
        assert_close(gt, r["grad_ctrl"], GTOL, "grad_ctrl vs repaired reference")


def _oracle_params(z):
    from oracle import macbf_oracle as O
    p_cbf = dict(zip(O.CBF_KEYS, [t.reshape(shp) for t, shp in zip(
        torch.split(torch.from_numpy(z["cbf_flat"]), [int(np.prod(sh)) for sh in O.CBF_SHAPES]), O.CBF_SHAPES)]))
    p_ctrl = dict(zip(O.CTRL_KEYS, [t.reshape(shp) for t, shp in zip(
        torch.split(torch.from_numpy(z["ctrl_flat"]), [int(np.prod(sh)) for sh in O.CTRL_SHAPES]), O.CTRL_SHAPES)]))
    return p_cbf, p_ctrl


@pytest.mark.parametrize("case", ["oracle_n96_obs.npz", "oracle_n8_small.npz", "oracle_n300_k16.npz"])
def test_cbf_backward_states_and_params(ops, case):
    """d(sum(h * w))/d(params, states) for a random cotangent w, against CPU autograd of the oracle."""
    from oracle import macbf_oracle as O
    z = load_npz(case)
    s, g, obs = _inputs(z)
    k = int(z["k"])
    p_cbf, _ = _oracle_params(z)
    gen = torch.Generator().manual_seed(1)
    w = torch.randn(z["h"][:, 0, :].shape, generator=gen)
    sc = torch.from_numpy(z["states"]).clone().requires_grad_(True)
    pc = {n: t.clone().requires_grad_(True) for n, t in p_cbf.items()}
    ob = torch.from_numpy(z["obstacles"]) if z["obstacles"].shape[0] else None
    h = O.cbf_forward(pc, sc, O.REPAIRED, ob, k)
    (h[:, 0, :] * w).sum().backward()
    ref_gp = torch.cat([pc[n].grad.reshape(-1) for n in O.CBF_KEYS]).numpy()
    idx = dev(z["idx"]).unsqueeze(0)
    gp, gs = ops.cbf_backward(s, idx, dev(z["cbf_flat"]), w.cuda().unsqueeze(0).contiguous(), obs, want_grad_states=True)
    assert_close(gp.cpu().numpy(), ref_gp, GTOL, "grad_params")
    assert_close(gs[0].cpu().numpy(), sc.grad.numpy(), GTOL, "grad_states")


@pytest.mark.parametrize("case", ["oracle_n96_obs.npz", "oracle_n8_small.npz", "oracle_n300_k16.npz"])
def test_controller_backward_states_goals_params(ops, case):
    from oracle import macbf_oracle as O
    z = load_npz(case)
    s, g, obs = _inputs(z)
    k = int(z["k"])
    _, p_ctrl = _oracle_params(z)
    gen = torch.Generator().manual_seed(2)
    w = torch.randn(z["u"].shape, generator=gen)
    wn = torch.randn(z["s_next"].shape, generator=gen)
    sc = torch.from_numpy(z["states"]).clone().requires_grad_(True)
    gc = torch.from_numpy(z["goals"]).clone().requires_grad_(True)
    pc = {n: t.clone().requires_grad_(True) for n, t in p_ctrl.items()}
    ob = torch.from_numpy(z["obstacles"]) if z["obstacles"].shape[0] else None
    u = O.controller_forward(pc, sc, gc, O.REPAIRED, ob, k)
    sn = O.euler_step(sc, u)
    ((u * w).sum() + (sn * wn).sum()).backward()
    ref_gp = torch.cat([pc[n].grad.reshape(-1) for n in O.CTRL_KEYS]).numpy()
    idx = dev(z["idx"]).unsqueeze(0)
    gp, gs, gg = ops.controller_backward(s, g, idx, dev(z["ctrl_flat"]), w.cuda().unsqueeze(0).contiguous(),
                                         wn.cuda().unsqueeze(0).contiguous(), obs, want_grad_states=True,
                                         want_grad_goals=True)
    assert_close(gp.cpu().numpy(), ref_gp, GTOL, "grad_params")
    assert_close(gs[0].cpu().numpy(), sc.grad.numpy(), GTOL, "grad_states")
    assert_close(gg[0].cpu().numpy(), gc.grad.numpy(), GTOL, "grad_goals")


@pytest.mark.parametrize("case", ["oracle_c1_n32.npz", "oracle_n300_k16.npz"])
def test_dropin_modules_autograd(ops, case):
    """The reference's own call sequence (train.py:88-105 for one state batch) through the drop-in
    core / cbf / controller modules: same losses and parameter gradients as the oracle."""
    import cbf as cbf_mod
    import controller as ctrl_mod
    import core
    from oracle import macbf_oracle as O
    z = load_npz(case)
    if int(z["k"]) != 12:
        old_k = core._cfg.TOP_K
        core._cfg.TOP_K = int(z["k"])
        cbf_mod.TOP_K = ctrl_mod.TOP_K = int(z["k"])
    try:
        net_h, net_u = cbf_mod.CBF(4).cuda(), ctrl_mod.Controller(4).cuda()
        p_cbf, p_ctrl = _oracle_params(z)
        net_h.load_state_dict({n: t for n, t in p_cbf.items()})
        net_u.load_state_dict({n: t for n, t in p_ctrl.items()})
        s, g = dev(z["states"]), dev(z["goals"])
        a = net_u(s, g)
        h = net_h(s)
        assert h.shape == z["h"].shape and a.shape == z["u"].shape
        lb = core.loss_barrier(h, s)
        ld = core.loss_derivatives(s, a, h, net_h)
        la = core.loss_actions(s, g, a)
        total = 10 * sum([2 * lb[0], lb[1], 2 * ld[0], ld[1], 0.01 * la])
        total.backward()
        got = [float(x) for x in (*lb, *ld, la)]
        for q in range(9):
            ref = float(z["scalars"][q])
            assert abs(got[q] - ref) <= 2 * RTOL * max(abs(ref), 1e-3), (q, got[q], ref)
        gcbf = torch.cat([p.grad.reshape(-1) for p in net_h._params()]).cpu().numpy()
        gctl = torch.cat([p.grad.reshape(-1) for p in net_u._params()]).cpu().numpy()
        assert_close(gcbf, z["grad_cbf"], GTOL, "grad_cbf")
        assert_close(gctl, z["grad_ctrl"], GTOL, "grad_ctrl")
    finally:
        if int(z["k"]) != 12:
            core._cfg.TOP_K = old_k
            cbf_mod.TOP_K = ctrl_mod.TOP_K = old_k


def test_env_sharded_gradient_equals_single_gpu(ops):
    """SURVEY 8e without a cluster: the env batch split into G logical shards, each with the
    global-mean seed, summed = the un-sharded gradient (what the NCCL all-reduce computes)."""
    from macbf_b200 import step
    from oracle import macbf_oracle as O
    B, N = 8, 200
    s, g = O.synthetic_swarm(B, N, seed=31)
    p_cbf, p_ctrl = O.init_params(4)
    cf, tf = dev(O.flatten_params(p_cbf, O.CBF_KEYS).numpy()), dev(O.flatten_params(p_ctrl, O.CTRL_KEYS).numpy())
    full = step.TrainStep(B, N, 0, 12, "cuda")
    _, g_full = full.forward_backward(s.cuda(), g.cuda(), cf, tf)
    g_full = g_full.clone()
    for G in (2, 4):
        acc = torch.zeros_like(g_full)
        for r in range(G):
            lo, hi = step.shard_envs(B, G, r)
            ts = step.TrainStep(hi - lo, N, 0, 12, "cuda", world_size=G)
            _, gr = ts.forward_backward(s[lo:hi].cuda().contiguous(), g[lo:hi].cuda().contiguous(), cf, tf,
                                        all_reduce=False)
            acc += gr
        assert rel_err(acc.cpu().numpy(), g_full.cpu().numpy()) <= 1e-5
    # and against the oracle's mean-over-envs objective
    ref = np.zeros(g_full.numel())
    for b in range(B):
        out = O.full_step(p_cbf, p_ctrl, s[b], g[b], with_grad=True)
        ref += np.concatenate([out["grad_cbf"].numpy(), out["grad_ctrl"].numpy()]) / B
    assert_close(g_full.cpu().numpy(), ref, GTOL, "batched gradient")


def test_dropin_module_sees_in_place_parameter_edits(ops):
    """Edits made through `p.data` do not bump the tensor version: the drop-in modules must not serve packed
    weights from before the edit (forward and backward)."""
    import cbf as cbf_mod
    import controller as ctrl_mod
    from oracle import macbf_oracle as O
    s, g = O.synthetic_swarm(1, 120, 77)
    s, g = s[0].cuda(), g[0].cuda()
    net_h, net_u = cbf_mod.CBF(4).cuda(), ctrl_mod.Controller(4).cuda()
    h0, u0 = net_h(s).detach().clone(), net_u(s, g).detach().clone()
    with torch.no_grad():
        for p in list(net_h.parameters()) + list(net_u.parameters()):
            p.data.mul_(0.5)
    h1, u1 = net_h(s), net_u(s, g)
    assert not torch.allclose(h0, h1) and not torch.allclose(u0, u1)
    p_cbf = {n: t.detach().cpu() for n, t in net_h.state_dict().items()}
    p_ctrl = {n: t.detach().cpu() for n, t in net_u.state_dict().items()}
    assert_close(h1.detach().cpu().numpy(), O.cbf_forward(p_cbf, s.cpu()).numpy(), RTOL, "h after p.data edit")
    assert_close(u1.detach().cpu().numpy(), O.controller_forward(p_ctrl, s.cpu(), g.cpu()).numpy(), RTOL, "u after edit")
