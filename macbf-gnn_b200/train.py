"""Joint training of the controller and the CBF certificate - the driver that sits on top of the
drop-in `core` / `cbf` / `controller` modules (same CLI and loop as the reference's train.py:17-109).

Differences from the reference's driver, all forced:
  * total loss = 10 * sum(terms)   (train.py:98 calls torch.add on a list and raises on any backend);
  * every tensor is created on the CUDA device (train.py:66 and the reference's modules create CPU
    tensors next to CUDA ones);
  * the per-iteration loss/accuracy lists keep floats, not autograd graphs (train.py:95-96 retain
    every iteration's graph forever);
  * --train_steps / --seed flags (defaults keep the reference behaviour: TRAIN_STEPS, unseeded).
The rollout is NOT detached, as in the reference (train.py:60-70): the backward pass goes through
every controller call of the trajectory, through the CUDA backward kernels.
"""
import argparse
import os

import numpy as np
import torch
import torch.optim as optim

import cbf
import controller
import core as _core
from config import *  # noqa: F401,F403
from core import *  # noqa: F401,F403


def parse_args(argv=None):
    parser = argparse.ArgumentParser()
    parser.add_argument('--num_agents', type=int, required=True)
    parser.add_argument('--model_path', type=str, default=None)
    parser.add_argument('--gpu', type=str, default='0')
    parser.add_argument('--train_steps', type=int, default=TRAIN_STEPS)
    parser.add_argument('--seed', type=int, default=None)
    parser.add_argument('--quiet', action='store_true')
    return parser.parse_args(argv)


def rollout_stepwise(policy, states, goals):
    """Up to INNER_LOOPS closed-loop Euler steps (train.py:59-81), one Python iteration and one host read of the stop
    test per step, as the reference does; returns the concatenated (states, goals, actions) of the visited time steps:
    T*N pseudo-agents in ONE graph (T2).  Kept as the cross-check of `rollout` (tests/test_gpu_train_driver.py)."""
    visited_s, visited_g, visited_a = [], [], []
    for _ in range(INNER_LOOPS):
        visited_s.append(states)
        visited_g.append(goals)
        actions = policy(states, goals)
        if np.random.uniform() < ADD_NOISE_PROB:
            actions = actions + torch.randn(actions.shape, device=actions.device) * NOISE_SCALE
        states = states + dynamics(states, actions) * TIME_STEP
        visited_a.append(actions)
        reached = torch.mean(torch.norm(states[:, :2] - goals, dim=1)) < DIST_MIN_CHECK
        if bool(reached):          # host sync per step, as in the reference (train.py:78)
            break
    return torch.cat(visited_s, dim=0), torch.cat(visited_g, dim=0), torch.cat(visited_a, dim=0)


def _draw_noise(n_steps, shape, device):
    """The reference's draws for `n_steps` visited steps (train.py:65-66): one np.random coin per step, one torch.randn
    per hit.  -> [n_steps, *shape] with zero rows where the coin said no, or None when no step drew noise."""
    hits = [np.random.uniform() < ADD_NOISE_PROB for _ in range(n_steps)]
    if not any(hits):
        return None
    noise = torch.zeros((n_steps,) + tuple(shape), device=device)
    for t, hit in enumerate(hits):
        if hit:
            noise[t] = torch.randn(shape, device=device) * NOISE_SCALE
    return noise


def rollout(policy, states, goals):
    """The same trajectory from ONE fused call (macbf_rollout_forward: the stop test runs on the device, the host reads
    the step count once) whose autograd is macbf_rollout_backward: BPTT through every controller call, as in the
    reference, without a Python graph of T nodes.  The random streams are consumed exactly as the step-by-step loop
    consumes them: noise for all INNER_LOOPS steps is drawn up front, and once the number of visited steps is known
    the generators are rewound and advanced by the visited steps only."""
    from macbf_b200 import functional as _F
    np_state = np.random.get_state()
    th_state = torch.cuda.get_rng_state(states.device)
    n = states.shape[0]
    noise = _draw_noise(INNER_LOOPS, (n, 2), states.device)
    obs = _core._obstacles_for(states.unsqueeze(0))
    ts, ta, steps = _F.RolloutFunction.apply(states.unsqueeze(0).contiguous(), goals.unsqueeze(0).contiguous(), obs,
                                             None if noise is None else noise.unsqueeze(1).contiguous(), INNER_LOOPS, TOP_K,
                                             DIST_MIN_CHECK, *policy._params())
    T = ts.shape[0]
    np.random.set_state(np_state)
    torch.cuda.set_rng_state(th_state, states.device)
    _draw_noise(T, (n, 2), states.device)
    return ts.reshape(T * n, 4), goals.repeat(T, 1), ta.reshape(T * n, 2)


def trajectory_losses(certificate, traj_s, traj_g, traj_a):
    """The five weighted loss terms and four accuracies of train.py:88-94."""
    h = certificate(traj_s)
    l_dang, l_safe, a_dang, a_safe = loss_barrier(h, traj_s)
    l_dang_d, l_safe_d, a_dang_d, a_safe_d = loss_derivatives(traj_s, traj_a, h, certificate)
    l_act = loss_actions(traj_s, traj_g, traj_a)
    terms = [2 * l_dang, l_safe, 2 * l_dang_d, l_safe_d, 0.01 * l_act]
    return terms, [a_dang, a_safe, a_dang_d, a_safe_d]


def main(argv=None):
    args = parse_args(argv)
    os.environ.setdefault("CUDA_VISIBLE_DEVICES", args.gpu)
    if not torch.cuda.is_available():
        raise RuntimeError("this driver runs the CUDA path only (no CPU fallback)")
    device = torch.device('cuda')
    if args.seed is not None:
        np.random.seed(args.seed)
        torch.manual_seed(args.seed)

    policy = controller.Controller(in_dim=4).to(device)
    certificate = cbf.CBF(in_dim=4).to(device)
    opt_policy = optim.Adam(policy.parameters(), lr=LEARNING_RATE, weight_decay=WEIGHT_DECAY)
    opt_certificate = optim.Adam(certificate.parameters(), lr=LEARNING_RATE, weight_decay=WEIGHT_DECAY)

    history = []
    steps = range(args.train_steps)
    if not args.quiet:
        try:
            from tqdm import tqdm
            steps = tqdm(steps)
        except ImportError:
            pass
    for _ in steps:
        s0, g0 = generate_data(args.num_agents, DIST_MIN_THRES)
        traj = rollout(policy, torch.from_numpy(s0).to(device), torch.from_numpy(g0).to(device))
        terms, accs = trajectory_losses(certificate, *traj)
        total = 10 * sum(terms)
        opt_policy.zero_grad()
        opt_certificate.zero_grad()
        total.backward()
        opt_policy.step()
        opt_certificate.step()
        history.append(([float(t) for t in terms], [float(a) for a in accs], float(total)))
    return history


if __name__ == '__main__':
    main()
