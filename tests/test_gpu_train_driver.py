"""GPU: the training driver on the drop-in modules (BASELINE config 1: train.py --num_agents 32).

One full iteration of the reference's loop (train.py:48-105: closed-loop rollout that is NOT
detached, one graph over the T*N trajectory states, five losses, backward through every controller
call) is replayed with the CPU oracle under torch autograd and compared: same rollout length, same
losses (1e-5), same parameter gradients after back-propagation through time (1e-4 scale-relative:
the chain is ~17 controller calls deep).
"""
import numpy as np
import pytest
import torch

from helpers import RTOL, assert_close

pytestmark = pytest.mark.gpu


def _oracle_iteration(p_cbf, p_ctrl, s0, g0):
    from oracle import macbf_oracle as O
    p_cbf = {n: t.clone().requires_grad_(True) for n, t in p_cbf.items()}
    p_ctrl = {n: t.clone().requires_grad_(True) for n, t in p_ctrl.items()}
    s, g = s0, g0
    S, G, A = [], [], []
    for _ in range(50):
        S.append(s); G.append(g)
        a = O.controller_forward(p_ctrl, s, g)
        s = O.euler_step(s, a)
        A.append(a)
        if torch.mean(torch.norm(s[:, :2] - g, dim=1)) < 0.07:
            break
    S, G, A = torch.cat(S), torch.cat(G), torch.cat(A)
    h = O.cbf_forward(p_cbf, S)
    lb = O.loss_barrier(h, S)
    ld, _ = O.loss_derivatives(p_cbf, S, A, h)
    la = O.loss_actions(S, G, A)
    total = O.total_loss(lb[0], lb[1], ld[0], ld[1], la)
    total.backward()
    gc = torch.cat([p_cbf[n].grad.reshape(-1) for n in O.CBF_KEYS])
    gt = torch.cat([p_ctrl[n].grad.reshape(-1) for n in O.CTRL_KEYS])
    return len(A) // s0.shape[0], float(total), gc.numpy(), gt.numpy()


def test_one_training_iteration_matches_oracle_bptt():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import cbf as cbf_mod
    import controller as ctrl_mod
    import core
    import train
    torch.manual_seed(0)
    policy = ctrl_mod.Controller(4).cuda()
    certificate = cbf_mod.CBF(4).cuda()
    np.random.seed(0)
    s0, g0 = core.generate_data(32, 0.07)
    # start with some velocity so that dangerous edges exist
    gen = torch.Generator().manual_seed(5)
    s0 = torch.from_numpy(s0)
    s0[:, 2:] = torch.randn(32, 2, generator=gen) * 0.3
    g0 = torch.from_numpy(g0)
    traj = train.rollout(policy, s0.cuda(), g0.cuda())
    terms, accs = train.trajectory_losses(certificate, *traj)
    total = 10 * sum(terms)
    total.backward()
    T = traj[0].shape[0] // 32

    p_cbf = {n: t.detach().cpu() for n, t in certificate.state_dict().items()}
    p_ctrl = {n: t.detach().cpu() for n, t in policy.state_dict().items()}
    T_ref, total_ref, gc_ref, gt_ref = _oracle_iteration(p_cbf, p_ctrl, s0, g0)
    assert T == T_ref
    assert abs(float(total) - total_ref) <= 5 * RTOL * abs(total_ref)
    gc = torch.cat([p.grad.reshape(-1) for p in certificate._params()]).cpu().numpy()
    gt = torch.cat([p.grad.reshape(-1) for p in policy._params()]).cpu().numpy()
    assert_close(gc, gc_ref, 1e-4, "grad_cbf (BPTT)")
    assert_close(gt, gt_ref, 1e-4, "grad_ctrl (BPTT)")


def test_fused_rollout_equals_the_step_by_step_loop():
    """train.rollout (one fused call + macbf_rollout_backward) against train.rollout_stepwise (the reference's Python
    loop over the per-step autograd Functions): same length, same trajectory, same gradients of a loss over it, and the
    same consumption of np.random (the draws that follow must agree)."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import controller as ctrl_mod
    import core
    import train
    torch.manual_seed(0)
    policy = ctrl_mod.Controller(4).cuda()
    np.random.seed(3)
    s0, g0 = core.generate_data(48, 0.07)
    s0 = torch.from_numpy(s0); g0 = torch.from_numpy(g0)
    s0[:, 2:] = torch.randn(48, 2, generator=torch.Generator().manual_seed(5)) * 0.2
    out = {}
    for name, fn in (("fused", train.rollout), ("stepwise", train.rollout_stepwise)):
        np.random.seed(11)
        policy.zero_grad()
        S, G, A = fn(policy, s0.cuda(), g0.cuda())
        w = torch.linspace(-1.0, 1.0, S.numel(), device="cuda").view_as(S)
        ((w * S).sum() + (A * A).sum()).backward()
        out[name] = (S.detach().cpu().numpy(), G.detach().cpu().numpy(), A.detach().cpu().numpy(),
                     torch.cat([p.grad.reshape(-1) for p in policy._params()]).cpu().numpy(), np.random.uniform())
    f, r = out["fused"], out["stepwise"]
    assert f[0].shape == r[0].shape and f[0].shape[0] > 48          # more than one step was visited
    assert_close(f[0], r[0], RTOL, "trajectory states")
    assert np.array_equal(f[1], r[1])
    assert_close(f[2], r[2], RTOL, "trajectory actions")
    assert_close(f[3], r[3], 1e-5, "grad_ctrl fused vs stepwise")
    assert f[4] == r[4], "np.random was consumed differently"


def test_driver_runs_and_updates_parameters():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import train
    hist = train.main(["--num_agents", "32", "--train_steps", "3", "--seed", "1", "--quiet"])
    assert len(hist) == 3
    for terms, accs, total in hist:
        assert np.isfinite(terms).all() and np.isfinite(total)
