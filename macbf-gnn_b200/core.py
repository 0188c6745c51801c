"""Drop-in for the reference's core.py: same function names, argument meaning and return shapes,
with the arithmetic of the hot path running in libmacbf_b200.so (CUDA, sm_100a).

REPAIRED semantics are the default because the reference's losses do not
execute as shipped; set ``core.SEMANTICS = LITERAL`` (or MACBF_SEMANTICS=literal) to get the
reference's forwards bug for bug (forward only).  Tensors must be on a CUDA device.
"""
import os as _os

import numpy as np
import torch

from config import *  # noqa: F401,F403  (re-exported, as the reference's core.py does: train.py:11-12)
import config as _cfg
from macbf_b200 import functional as _F
from macbf_b200 import ops as _ops

LITERAL = _ops.LITERAL
REPAIRED = _ops.REPAIRED
SEMANTICS = LITERAL if _os.environ.get("MACBF_SEMANTICS", "repaired").lower() == "literal" else REPAIRED
# optional static obstacle points (M,2) / (B,M,2) used as extra candidate columns (extension)
OBSTACLES = None


# ------------------------------------------------------------------ scenario generation (host)
def generate_obstacle_circle(center, radius, num=12):
    """`num` points on a circle (core.py:7-12), float64 (num,2)."""
    ang = np.linspace(0, np.pi * 2, num=num, endpoint=False)
    return np.asarray(center) + radius * np.stack([np.cos(ang), np.sin(ang)], axis=1)


def generate_obstacle_rectangle(center, sides, num=12):
    """`num` points on a rectangle's perimeter, clockwise from the top-left corner (core.py:15-42)."""
    w, hgt = sides
    n_top = int(num // 2 * w / (w + hgt))
    n_right = num // 2 - n_top
    n_bottom = n_top
    n_left = num - n_top - n_right - n_bottom
    top = np.stack([np.linspace(-w / 2, w / 2, n_top, endpoint=False), np.full(n_top, hgt / 2)], axis=1)
    right = np.stack([np.full(n_right, w / 2), np.linspace(hgt / 2, -hgt / 2, n_right, endpoint=False)], axis=1)
    bottom = np.stack([np.linspace(w / 2, -w / 2, n_bottom, endpoint=False), np.full(n_bottom, -hgt / 2)], axis=1)
    left = np.stack([np.full(n_left, -w / 2), np.linspace(-hgt / 2, hgt / 2, n_left, endpoint=False)], axis=1)
    return np.concatenate([top, right, bottom, left], axis=0) + np.asarray(center)


def generate_obstacles_cuda(specs, num=12, device="cuda"):
    """generate_obstacle_circle / generate_obstacle_rectangle for a list of obstacles on the device
    (macbf_generate_obstacles): specs = [("circle", center, radius) | ("rectangle", center, (a, b)), ...]
    -> CUDA float64 tensor (len(specs) * num, 2), the concatenation of the reference's point sets."""
    return _ops.generate_obstacles(specs, num, device)


def _rejection_fill(out, draw, min_sep):
    """Fill `out` row by row; a draw is rejected when it is within `min_sep` of ANY row of `out`,
    including rows still at their initial zero (the reference's behaviour, core.py:51-56)."""
    i = 0
    while i < out.shape[0]:
        cand = draw(i)
        if np.min(np.linalg.norm(out - cand, axis=1)) > min_sep:
            out[i] = cand
            i += 1


def generate_data(num_agents, dist_min_thres):
    """Random starts in a sqrt(max(1,N/8)) square and goals within +-0.5 of them, all pairwise
    farther than dist_min_thres; consumes np.random exactly like core.py:45-71.
    Returns (states (N,4) float32 with zero velocity, goals (N,2) float32)."""
    side = np.sqrt(max(1.0, num_agents / 8.0))
    pos = np.zeros((num_agents, 2), dtype=np.float32)
    goals = np.zeros((num_agents, 2), dtype=np.float32)
    _rejection_fill(pos, lambda i: np.random.uniform(size=(2,)) * side, dist_min_thres)
    _rejection_fill(goals, lambda i: np.random.uniform(-0.5, 0.5, size=(2,)) + pos[i], dist_min_thres)
    return np.concatenate([pos, np.zeros_like(pos)], axis=1), goals


def generate_data_cuda(num_agents, dist_min_thres, seed=0, num_envs=None, device="cuda"):
    """generate_data on the device (macbf_generate_scenario): the same dart-throwing process on a counter-based
    stream keyed by (seed, env) instead of np.random.  -> CUDA tensors (states (N,4), goals (N,2)), or batched
    (B,N,4), (B,N,2) when num_envs is given."""
    s, g = _ops.generate_scenario(seed, 1 if num_envs is None else int(num_envs), int(num_agents), dist_min_thres, device)
    return (s[0], g[0]) if num_envs is None else (s, g)


# ------------------------------------------------------------------ hot path
def _obstacles_for(states_b):
    if OBSTACLES is None:
        return None
    o = torch.as_tensor(OBSTACLES, dtype=torch.float32, device=states_b.device)
    if o.dim() == 2:
        o = o.unsqueeze(0).expand(states_b.shape[0], -1, -1)
    return o.contiguous()


def neighbor_graph(states, k=None):
    """idx [B,N,K_eff] int32 for states [N,4] or [B,N,4] (the index half of core.py:234-250)."""
    sb, _ = _F._batched(states.detach(), 4)
    return _ops.neighbor_topk(sb, _cfg.TOP_K if k is None else k, _obstacles_for(sb))


def dynamics(states, actions):
    """Double integrator ds/dt = [v, a] (core.py:74-86)."""
    return torch.cat([states[..., 2:], actions], dim=-1)


def remove_distant_agents(x, k):
    """(N, C, F) pair tensor -> ((N, k, F), indices (N*k, 2) int64), k nearest by the key of
    core.py:238 on x[:, :, :2].  With <= k candidates returns (x, None) (repair F3)."""
    n, c, _ = x.shape
    if c <= k:
        return x, None
    idx = _ops.dense_topk(x, k).long()
    rows = torch.arange(n, device=x.device).unsqueeze(1).expand_as(idx)
    gathered = x[rows, idx]
    return gathered, torch.stack([rows.reshape(-1), idx.reshape(-1)], dim=1)


def ttc_dangerous_mask(s):
    """bool (N,K,1) (or (B,N,K,1)): within DIST_MIN_THRES or colliding within TIME_TO_COLLISION
    (core.py:187-209), on the top-k edges."""
    sb, squeeze = _F._batched(s.detach(), 4)
    obs = _obstacles_for(sb)
    idx = _ops.neighbor_topk(sb, _cfg.TOP_K, obs)
    m = _ops.ttc_mask(sb, idx, obs, _cfg.DIST_MIN_THRES, _cfg.TIME_TO_COLLISION).bool().unsqueeze(-1)
    return m[0] if squeeze else m


def ttc_dangerous_mask_np(s):
    """Dense (N,N,1) evaluation-time check with the *_CHECK constants (core.py:212-231).  A numpy array is
    evaluated on the host exactly as the reference does; a CUDA tensor (N,4) / (B,N,4) goes through
    macbf_ttc_dense (same bits) and returns a bool tensor (N,N,1) / (B,N,N,1)."""
    if isinstance(s, torch.Tensor):
        sb, squeeze = _F._batched(s.detach(), 4)
        _, m = _ops.ttc_dense(sb, _cfg.DIST_MIN_CHECK, _cfg.TIME_TO_COLLISION_CHECK, want_counts=False)
        m = m.bool().unsqueeze(-1)
        return m[0] if squeeze else m
    d = s[:, None, :] - s[None, :, :]
    eye = np.eye(s.shape[0])[:, :, None]
    x, y, vx, vy = d[:, :, 0:1] + eye, d[:, :, 1:2] + eye, d[:, :, 2:3], d[:, :, 3:4]
    alpha = vx ** 2 + vy ** 2
    beta = 2 * (x * vx + y * vy)
    gamma = x ** 2 + y ** 2 - _cfg.DIST_MIN_CHECK ** 2
    disc = beta ** 2 - 4 * alpha * gamma
    two_positive_roots = (disc > 0) & (gamma > 0) & (beta < 0)
    hits_soon = (-beta - 2 * alpha * _cfg.TIME_TO_COLLISION_CHECK < 0) | \
        ((beta + 2 * alpha * _cfg.TIME_TO_COLLISION_CHECK) ** 2 < disc)
    return (gamma < 0) | (two_positive_roots & hits_soon)


def _h_batched(h):
    """(N,1,K) / (B,N,1,K) / (N,K) / (B,N,K) -> [B,N,K]."""
    if h.dim() == 4:
        return h.squeeze(2)
    if h.dim() == 3 and h.shape[1] == 1:
        return h.squeeze(1).unsqueeze(0)
    if h.dim() == 2:
        return h.unsqueeze(0)
    return h


def _mean_over_envs(scalars, cols):
    # per-env losses averaged over the env batch (B = 1 for the reference's call shapes);
    # accuracies of envs whose set is empty are -1 as in core.py:124-127
    return tuple(scalars[:, c].mean() for c in cols)


def loss_barrier(h, states):
    """(loss_dang, loss_safe, acc_dang, acc_safe) - core.py:89-129."""
    sb, _ = _F._batched(states, 4)
    hb = _h_batched(h)
    obs = _obstacles_for(sb)
    idx = _ops.neighbor_topk(sb.detach(), _cfg.TOP_K, obs)
    scalars, _ = _F.HingeLossFunction.apply(hb.contiguous(), None, None, sb.detach(), None, obs, idx)
    return _mean_over_envs(scalars, (0, 1, 2, 3))


def loss_derivatives(states, actions, h, cbf):
    """(loss_dang_deriv, loss_safe_deriv, acc_dang_deriv, acc_safe_deriv) - core.py:132-171;
    h_next is the CBF re-evaluated on the Euler-stepped states with a fresh neighbour graph."""
    states_next = states + dynamics(states, actions) * _cfg.TIME_STEP
    h_next = cbf.forward(states_next)
    sb, _ = _F._batched(states, 4)
    obs = _obstacles_for(sb)
    idx = _ops.neighbor_topk(sb.detach(), _cfg.TOP_K, obs)
    scalars, _ = _F.HingeLossFunction.apply(_h_batched(h).contiguous(), _h_batched(h_next).contiguous(), None,
                                            sb.detach(), None, obs, idx)
    return _mean_over_envs(scalars, (4, 5, 6, 7))


def loss_actions(s, g, a):
    """mean | |a|^2 - |a_ref|^2 | against the LQR-like reference a_ref = -(p-g) - sqrt(3) v (core.py:174-184)."""
    sb, _ = _F._batched(s, 4)
    gb, _ = _F._batched(g, 2)
    ab, _ = _F._batched(a, 2)
    k_eff = _ops.neighbor_count(sb.shape[1], 0, 1)
    idx = torch.zeros((sb.shape[0], sb.shape[1], k_eff), dtype=torch.int32, device=sb.device)
    idx += torch.arange(sb.shape[1], dtype=torch.int32, device=sb.device).view(1, -1, 1)
    dummy_h = torch.zeros((sb.shape[0], sb.shape[1], k_eff), dtype=torch.float32, device=sb.device)
    scalars, _ = _F.HingeLossFunction.apply(dummy_h, None, ab.contiguous(), sb.contiguous(), gb.detach(), None, idx)
    return scalars[:, 8].mean()
