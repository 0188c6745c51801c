"""GPU parity: macbf_rollout_forward (closed-loop rollout of train.py:59-81, stop test on the device)
against the oracle's controller + Euler step driven by the same loop on the host."""
import numpy as np
import pytest
import torch

from helpers import RTOL, assert_close, dev

pytestmark = pytest.mark.gpu

DIST_STOP = 0.07   # config.py DIST_MIN_CHECK


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from macbf_b200 import ops as _ops
    return _ops


def _oracle_rollout(O, p_ctrl, s0, g, obs, T, noise=None):
    """train.py:59-81 for ONE env: lists of visited states / actions, and the final state."""
    s = s0.clone()
    visited, actions = [], []
    for t in range(T):
        visited.append(s)
        a = O.controller_forward(p_ctrl, s, g, O.REPAIRED, obs)
        if noise is not None:
            a = a + noise[t]
        s = O.euler_step(s, a)
        actions.append(a)
        if torch.mean(torch.norm(s[:, :2] - g, dim=1)) < DIST_STOP:
            break
    return visited, actions, s


def _swarm(O, N, seed):
    """3 envs: env 0 starts (almost) at its goals and stops after the first step, env 1 is a generic swarm
    that keeps running, env 2 is close enough to stop after a few steps."""
    s, g = O.synthetic_swarm(3, N, seed)
    s[0, :, 2:] = 0.0
    g[0] = s[0, :, :2] + 0.02
    s[2, :, 2:] = 0.05
    g[2] = s[2, :, :2] + 0.055       # stops after its 2nd step with a 6e-3 margin to the threshold
    return s, g


@pytest.mark.parametrize("N,n_obs,with_noise", [(32, 0, False), (300, 2, False), (64, 0, True)])
def test_rollout_matches_oracle_loop(ops, N, n_obs, with_noise):
    import oracle.macbf_oracle as O
    T = 6
    _, p_ctrl = O.init_params(5)
    flat = dev(O.flatten_params(p_ctrl, O.CTRL_KEYS))
    s, g = _swarm(O, N, 11)
    obs = O.synthetic_obstacles(3, N, n_obs, 0, 4) if n_obs else None
    noise = None
    if with_noise:
        gen = torch.Generator().manual_seed(3)
        noise = torch.randn(T, 3, N, 2, generator=gen) * 0.3
        noise[1] = 0.0                       # a step whose coin said "no noise"
        noise[:, 0] = 0.0                    # env 0 must still stop at once
    ts, ta, steps = ops.rollout_forward(dev(s), dev(g), flat, T, obstacles=None if obs is None else dev(obs),
                                        noise=None if noise is None else dev(noise))
    ts, ta, steps = ts.cpu(), ta.cpu(), steps.cpu().numpy()
    assert torch.equal(ts[0], s)
    ran = []
    for b in range(3):
        ob = None if obs is None else obs[b]
        vis, act, _ = _oracle_rollout(O, p_ctrl, s[b], g[b], ob, T, None if noise is None else noise[:, b])
        ran.append(len(vis))
        assert steps[b] == len(vis), f"env {b}: steps_run {steps[b]} vs oracle {len(vis)}"
        for t in range(len(vis)):
            # free-running comparison (errors compound over the steps: 10x the one-step tolerance)
            assert_close(ts[t, b].numpy(), vis[t].numpy(), 10 * RTOL, f"s[{t}] env {b}")
            assert_close(ta[t, b].numpy(), act[t].numpy(), 10 * RTOL, f"a[{t}] env {b}")
            # one-step comparison from the GPU's own state at the one-step tolerance
            a1 = O.controller_forward(p_ctrl, ts[t, b], g[b], O.REPAIRED, ob)
            if noise is not None:
                a1 = a1 + noise[t, b]
            assert_close(ta[t, b].numpy(), a1.numpy(), RTOL, f"one-step a[{t}] env {b}")
            assert_close(ts[t + 1, b].numpy(), O.euler_step(ts[t, b], a1).numpy(), RTOL, f"one-step s[{t + 1}] env {b}")
    assert ran[0] == 1 and ran[1] == T, ran
    if not with_noise:
        assert 1 < ran[2] < T, ran                            # the three regimes were exercised


def test_rollout_all_stopped_skips_remaining_steps(ops):
    """Once every env has stopped, the remaining steps' kernels return at once: their trajectory
    slots stay untouched (zero) and steps_run says how many were recorded."""
    import oracle.macbf_oracle as O
    _, p_ctrl = O.init_params(5)
    flat = dev(O.flatten_params(p_ctrl, O.CTRL_KEYS))
    N, T = 2048, 8            # large enough for the tensor-core controller and the grid neighbour search
    s, g = O.synthetic_swarm(2, N, 2)
    s[:, :, 2:] = 0.0
    g = s[:, :, :2] + 0.01
    ts, ta, steps = ops.rollout_forward(dev(s), dev(g), flat, T)
    assert steps.cpu().tolist() == [1, 1]
    assert float(ts[1].abs().sum()) > 0 and float(ta[0].abs().sum()) > 0
    assert float(ts[2:].abs().sum()) == 0.0 and float(ta[1:].abs().sum()) == 0.0
    # matches the per-step API
    idx = ops.neighbor_topk(dev(s), 12)
    u, sn = ops.controller_forward(dev(s), dev(g), idx, flat)[:2]
    assert torch.equal(ta[0], u) and torch.equal(ts[1], sn)


def test_rollout_rejects_small_workspace(ops):
    from macbf_b200 import lib
    L = lib.load()
    s = torch.zeros(1, 64, 4, device="cuda"); g = torch.zeros(1, 64, 2, device="cuda")
    flat = torch.zeros(34052, device="cuda")
    ts = torch.zeros(3, 1, 64, 4, device="cuda"); ta = torch.zeros(2, 1, 64, 2, device="cuda")
    st = torch.zeros(1, dtype=torch.int32, device="cuda"); ws = torch.zeros(64, dtype=torch.uint8, device="cuda")
    rc = L.macbf_rollout_forward(s.data_ptr(), g.data_ptr(), None, flat.data_ptr(), None, 1, 64, 0, 12, 2, 1, 1.0, 0.05,
                                 0.07, ts.data_ptr(), ta.data_ptr(), st.data_ptr(), ws.data_ptr(), 64, None)
    assert rc == 1 and b"workspace too small" in L.macbf_last_error()


@pytest.mark.parametrize("N,n_obs,with_noise", [(32, 0, False), (200, 2, True)])
def test_rollout_backward_matches_oracle_bptt(ops, N, n_obs, with_noise):
    """macbf_rollout_backward (BPTT through the saved trajectory, train.py:59-81 is not detached) against torch
    autograd of the oracle's loop, for envs that stop after 1, a few and all T steps; the loss touches every visited
    state and action of every env.  Gradients on slots beyond an env's own steps_run must not count."""
    import oracle.macbf_oracle as O
    from macbf_b200 import functional as F
    T = 6
    _, p_ctrl = O.init_params(5)
    s, g = _swarm(O, N, 11)
    obs = O.synthetic_obstacles(3, N, n_obs, 0, 4) if n_obs else None
    noise = None
    if with_noise:
        gen = torch.Generator().manual_seed(3)
        noise = torch.randn(T, 3, N, 2, generator=gen) * 0.05
        noise[1] = 0.0
        noise[:, 0] = 0.0
    gen = torch.Generator().manual_seed(9)
    cs = torch.randn(T, 3, N, 4, generator=gen)         # L = sum_live cs . s_t + 0.5 |a_t|^2 + ca . a_t
    ca = torch.randn(T, 3, N, 2, generator=gen)

    params = [dev(p_ctrl[k]).requires_grad_(True) for k in O.CTRL_KEYS]
    s_d, g_d = dev(s).requires_grad_(True), dev(g).requires_grad_(True)
    ts, ta, steps = F.RolloutFunction.apply(s_d, g_d, None if obs is None else dev(obs), None if noise is None else dev(noise),
                                            T, 12, DIST_STOP, *params)
    Tm = ts.shape[0]
    # NO mask here: dead slots (an env's continued simulation) get gradients from this loss, the Function drops them
    loss = (dev(cs)[:Tm] * ts).sum() + (0.5 * ta * ta + dev(ca)[:Tm] * ta).sum()
    loss.backward()
    steps = steps.cpu().numpy()

    ref_p = {k: p_ctrl[k].clone().requires_grad_(True) for k in O.CTRL_KEYS}
    s_r, g_r = s.clone().requires_grad_(True), g.clone().requires_grad_(True)
    total = 0.0
    for b in range(3):
        ob = None if obs is None else obs[b]
        vis, act, _ = _oracle_rollout(O, ref_p, s_r[b], g_r[b], ob, T, None if noise is None else noise[:, b])
        assert len(vis) == steps[b]
        for t in range(len(vis)):
            total = total + (cs[t, b] * vis[t]).sum() + (0.5 * act[t] * act[t] + ca[t, b] * act[t]).sum()
    total.backward()
    assert steps.tolist()[0] == 1 and steps[1] == T
    gp = torch.cat([p.grad.reshape(-1) for p in params]).cpu().numpy()
    gp_ref = torch.cat([ref_p[k].grad.reshape(-1) for k in O.CTRL_KEYS]).numpy()
    assert_close(gp, gp_ref, 1e-4, "grad_ctrl (rollout BPTT)")
    assert_close(s_d.grad.cpu().numpy(), s_r.grad.numpy(), 1e-4, "d/ds_0 (rollout BPTT)")
    assert_close(g_d.grad.cpu().numpy(), g_r.grad.numpy(), 1e-4, "d/dgoals (rollout BPTT)")


def test_rollout_backward_rejects_small_scratch(ops):
    from macbf_b200 import lib
    L = lib.load()
    ts = torch.zeros(3, 1, 64, 4, device="cuda"); g = torch.zeros(1, 64, 2, device="cuda")
    flat = torch.zeros(34052, device="cuda"); ga = torch.zeros(2, 1, 64, 2, device="cuda")
    gp = torch.zeros(34052, device="cuda"); ws = torch.zeros(64, dtype=torch.uint8, device="cuda")
    rc = L.macbf_rollout_backward(ts.data_ptr(), g.data_ptr(), None, flat.data_ptr(), None, ga.data_ptr(), 1, 64, 0, 12, 2, 1.0,
                                  0.05, gp.data_ptr(), None, None, ws.data_ptr(), 64, None)
    assert rc == 1 and b"scratch too small" in L.macbf_last_error()
    assert L.macbf_rollout_backward_scratch_bytes(1, 64, 0, 12) > 64
