"""CPU oracle for the macbf-gnn per-step safety-certificate path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path (``macbf-gnn_b200/``) may import
this file; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs use it, and only as the checker / the timed CPU arm.

What it is
----------
A torch-CPU restatement of the reference's algorithm (``/root/reference``; 5 Python
files, no tests, no golden vectors).  Two semantics are provided:

* ``LITERAL``  - bug-for-bug what the reference's two forwards compute today
  (``cbf.py:21-45``, ``controller.py:31-63``): channels-last -> channels-first done by
  ``reshape`` (``cbf.py:40``, ``controller.py:43``) and arg-max *indices* pooled
  (``controller.py:46``).  Only the forwards exist under LITERAL (every loss crashes).
* ``REPAIRED`` - the enumerated repairs of SURVEY Appendix A (F1, F2, F3, L1-L9):
  a true transpose, max *values*, and loss functions that run.

Parity pin
----------
The reference ships no tests or fixtures.  The pin is the reference itself: the LITERAL
functions below are checked bit-for-bit against the unmodified reference modules
imported through ``oracle/ref_loader.py`` (``tests/golden/make_golden.py`` does this and
commits the resulting vectors; ``tests/test_oracle_golden.py`` re-checks the restatement
against those vectors on every run).  REPAIRED differs from LITERAL only by the repairs
listed above; the loss functions have no executable counterpart in the reference
("parity unpinned" for those rows: they follow ``core.py:89-209`` line by line).

Extensions beyond the reference (defined in SURVEY section 8 A0/(b))
--------------------------------------------------------------------
* ``obstacles`` (M,2): extra *candidate columns* ``[ox, oy, 0, 0]`` with ``eye = 0``;
  never query rows.  Neighbour indices then live in ``[0, N+M)``.
* ``top_k`` / radius / thresholds are arguments (defaults = ``config.py``).
* batched helpers apply the per-env function independently (``*_batched``).
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

LITERAL = 0
REPAIRED = 1

# config.py:1-14 (values the path consumes)
TIME_STEP = 1e-1
ALPHA_CBF = 1.0
DIST_MIN_THRES = 0.07
OBS_RADIUS = 1.0
TOP_K = 12
TIME_TO_COLLISION = 2.0

CBF_KEYS = [f"cbf_net.{i}.{p}" for i in (0, 2, 4, 6) for p in ("weight", "bias")]
CTRL_KEYS = [f"controller_centr_net.{i}.{p}" for i in (0, 2) for p in ("weight", "bias")] + [
    f"controller_dec_net.{i}.{p}" for i in (0, 2, 4, 6) for p in ("weight", "bias")
]
CBF_SHAPES = [(64, 6, 1), (64,), (128, 64, 1), (128,), (64, 128, 1), (64,), (1, 64, 1), (1,)]
CTRL_SHAPES = [(64, 5, 1), (64,), (128, 64, 1), (128,), (64, 132), (64,), (128, 64), (128,),
               (64, 128), (64,), (4, 64), (4,)]


# ----------------------------------------------------------------------------------------
# parameters
# ----------------------------------------------------------------------------------------
def init_params(seed: int, dtype=torch.float32) -> Tuple[Dict[str, torch.Tensor], Dict[str, torch.Tensor]]:
    """Random parameters with the reference's shapes and state_dict keys
    (cbf.py:13-19, controller.py:16-29).  Uniform(-1/sqrt(fan_in), 1/sqrt(fan_in))."""
    g = torch.Generator().manual_seed(seed)

    def make(keys, shapes):
        out = {}
        for k, shp in zip(keys, shapes):
            if k.endswith("weight"):
                fan_in = shp[1]
                last_fan = fan_in
            else:
                fan_in = last_fan
            bound = 1.0 / math.sqrt(fan_in)
            out[k] = ((torch.rand(shp, generator=g, dtype=torch.float64) * 2 - 1) * bound).to(dtype)
        return out

    last_fan = 1
    return make(CBF_KEYS, CBF_SHAPES), make(CTRL_KEYS, CTRL_SHAPES)


def flatten_params(p: Dict[str, torch.Tensor], keys) -> torch.Tensor:
    return torch.cat([p[k].reshape(-1) for k in keys])


# ----------------------------------------------------------------------------------------
# neighbour selection  (core.py:234-250)
# ----------------------------------------------------------------------------------------
def candidate_states(states: torch.Tensor, obstacles: Optional[torch.Tensor]) -> torch.Tensor:
    """Candidate columns: the agents themselves, then obstacle points as [ox, oy, 0, 0]."""
    if obstacles is None or obstacles.shape[0] == 0:
        return states
    obs = torch.zeros(obstacles.shape[0], 4, dtype=states.dtype)
    obs[:, :2] = obstacles.to(states.dtype)
    return torch.cat([states, obs], dim=0)


def pair_tensor(states: torch.Tensor, obstacles: Optional[torch.Tensor], rows: Optional[torch.Tensor] = None
                ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Dense (N, N+M, 4) relative states and the (N, N+M) identity indicator
    (cbf.py:23,28 / controller.py:33-34 / core.py:188-190).  rows (R,) int64: only these query rows (the row-chunked
    form for N where the dense tensor does not fit; every row is independent of the others)."""
    cand = candidate_states(states, obstacles)
    q = states if rows is None else states[rows]
    diff = q.unsqueeze(1) - cand.unsqueeze(0)
    n, c = q.shape[0], cand.shape[0]
    eye = torch.zeros(n, c, dtype=states.dtype)
    eye[torch.arange(n), torch.arange(n) if rows is None else rows] = 1.0
    return diff, eye


def knn_keys(xy_diff: torch.Tensor) -> torch.Tensor:
    """core.py:238 - sqrt(sum(square(dxy) + 1e-6))."""
    return torch.sqrt(torch.sum(torch.square(xy_diff) + 1e-6, dim=2))


def remove_distant_agents(x: torch.Tensor, k: int):
    """core.py:234-250.  (N, C, F) -> ((N, k, F), idx (N, k) int64).
    N <= k (rows here means candidate count C <= k): no selection, K_eff = C, index order
    (repair F3: the reference returns the bare tensor there)."""
    n, c, _ = x.shape
    if c <= k:
        return x, None
    key = knn_keys(x[:, :, :2])
    idx = torch.topk(-key, k)[1]                      # core.py:239
    rows = torch.arange(n).unsqueeze(1).expand_as(idx)
    return x[rows, idx], idx                          # core.py:240-248


def neighbor_indices(states: torch.Tensor, k: int, obstacles: Optional[torch.Tensor] = None,
                     row_chunk: int = 0) -> torch.Tensor:
    """Indices only, (N, K_eff) int64; row-chunked so dense N x N never has to fit
    (identical result: top-k is independent per row)."""
    cand = candidate_states(states, obstacles)
    n, c = states.shape[0], cand.shape[0]
    if c <= k:
        return torch.arange(c).unsqueeze(0).expand(n, c).contiguous()
    if row_chunk <= 0:
        row_chunk = n
    out = []
    for r0 in range(0, n, row_chunk):
        d = states[r0:r0 + row_chunk, None, :2] - cand[None, :, :2]
        out.append(torch.topk(-knn_keys(d), k)[1])
    return torch.cat(out, dim=0)


def neighbor_keys_sorted(states: torch.Tensor, k: int, obstacles: Optional[torch.Tensor] = None):
    """Sorted keys of the best k+1 candidates per row (for the tie/near-tie classifier,
    SURVEY Appendix B).  Returns (keys (N, min(k+1,C)), idx)."""
    cand = candidate_states(states, obstacles)
    d = states[:, None, :2] - cand[None, :, :2]
    key = knn_keys(d)
    kk = min(k + 1, cand.shape[0])
    v, i = torch.topk(-key, kk)
    return -v, i


def decidable_rows(states: torch.Tensor, k: int, obstacles: Optional[torch.Tensor] = None,
                   min_gap_ulp: int = 2) -> torch.Tensor:
    """A row is decidable when consecutive sorted keys among its best k+1 differ by at
    least ``min_gap_ulp`` ulp (Appendix B parity contract)."""
    keys, _ = neighbor_keys_sorted(states, k, obstacles)
    bits = keys.contiguous().view(torch.int32).to(torch.int64)
    gap = bits[:, 1:] - bits[:, :-1]
    return (gap >= min_gap_ulp).all(dim=1)


# ----------------------------------------------------------------------------------------
# edge features shared by CBF / controller / TTC
# ----------------------------------------------------------------------------------------
def _gather_edges(states, obstacles, k, extra_cbf_feature: bool, indices: Optional[torch.Tensor] = None,
                  rows: Optional[torch.Tensor] = None):
    """indices (N, K) int64: gather these candidate columns instead of a fresh top-k (the `indices=` argument of the
    code the reference derives from; the reference recomputes, core.py:137) - an opt-in extension.
    rows: restrict the query rows (see pair_tensor)."""
    diff, eye = pair_tensor(states, obstacles, rows)
    feats = [diff, eye.unsqueeze(2)]
    if extra_cbf_feature:
        d = torch.sqrt(torch.sum(torch.square(diff[:, :, :2]) + 1e-4, dim=2))   # cbf.py:24-25
        feats.append((d - DIST_MIN_THRES).unsqueeze(2))                         # cbf.py:29
    x = torch.cat(feats, dim=2)
    if indices is not None:
        r = torch.arange(x.shape[0]).unsqueeze(1).expand_as(indices)
        return x[r, indices], indices
    if x.shape[1] <= k:            # the reference tests the row count (core.py:236); with a row subset the candidate
        return x, None             # count is what decides
    key = knn_keys(x[:, :, :2])
    idx = torch.topk(-key, k)[1]                      # core.py:239
    r = torch.arange(x.shape[0]).unsqueeze(1).expand_as(idx)
    return x[r, idx], idx                             # core.py:240-248


def _to_channels_first(x: torch.Tensor, semantics: int) -> torch.Tensor:
    n, k, c = x.shape
    if semantics == LITERAL:
        return x.reshape(n, c, k)          # cbf.py:40 / controller.py:43 (memory reinterpretation)
    return x.permute(0, 2, 1)              # repair F1


# ----------------------------------------------------------------------------------------
# CBF  (cbf.py:21-45)
# ----------------------------------------------------------------------------------------
def cbf_forward(p: Dict[str, torch.Tensor], states: torch.Tensor, semantics: int = REPAIRED,
                obstacles: Optional[torch.Tensor] = None, k: int = TOP_K,
                obs_radius: float = OBS_RADIUS, indices: Optional[torch.Tensor] = None,
                rows: Optional[torch.Tensor] = None) -> torch.Tensor:
    """h (N, 1, K_eff)  [(R, 1, K_eff) for a row subset]."""
    x, _ = _gather_edges(states, obstacles, k, True, indices, rows)
    dist = torch.sqrt(torch.sum(torch.square(x[:, :, :2]) + 1e-4, dim=2, keepdim=True))  # cbf.py:34-35
    mask = torch.less_equal(dist, obs_radius).type_as(x)
    mask = mask.reshape(-1, mask.shape[2], mask.shape[1])                                # cbf.py:37
    y = _to_channels_first(x, semantics)
    y = F.relu(F.conv1d(y, p["cbf_net.0.weight"], p["cbf_net.0.bias"]))
    y = F.relu(F.conv1d(y, p["cbf_net.2.weight"], p["cbf_net.2.bias"]))
    y = F.relu(F.conv1d(y, p["cbf_net.4.weight"], p["cbf_net.4.bias"]))
    y = F.conv1d(y, p["cbf_net.6.weight"], p["cbf_net.6.bias"])
    return y * mask                                                                      # cbf.py:43


# ----------------------------------------------------------------------------------------
# Controller  (controller.py:31-63)
# ----------------------------------------------------------------------------------------
def controller_edge_values(p, states, semantics=REPAIRED, obstacles=None, k=TOP_K, obs_radius=OBS_RADIUS, rows=None):
    """Edge net output times the sensing mask, (N, 128, K_eff): what controller.py:46 pools over."""
    x, _ = _gather_edges(states, obstacles, k, False, None, rows)
    dist = torch.norm(x[:, :, :2], dim=2, keepdim=True)                                  # controller.py:38
    mask = torch.less(dist, obs_radius).type_as(x)
    mask = mask.reshape(-1, mask.shape[2], mask.shape[1])
    y = _to_channels_first(x, semantics)
    y = F.relu(F.conv1d(y, p["controller_centr_net.0.weight"], p["controller_centr_net.0.bias"]))
    y = F.relu(F.conv1d(y, p["controller_centr_net.2.weight"], p["controller_centr_net.2.bias"]))
    return y * mask


def controller_pool(p, states, semantics=REPAIRED, obstacles=None, k=TOP_K, obs_radius=OBS_RADIUS, rows=None):
    """Edge net + masked pooling -> (N, 128).  LITERAL pools arg-max indices (int64)."""
    pooled = torch.max(controller_edge_values(p, states, semantics, obstacles, k, obs_radius, rows), dim=2)  # controller.py:46
    return pooled[1] if semantics == LITERAL else pooled[0]


def controller_forward(p, states, goals, semantics=REPAIRED, obstacles=None, k=TOP_K,
                       obs_radius=OBS_RADIUS, rows=None) -> torch.Tensor:
    """a (N, 2)  [(R, 2) for a row subset]."""
    z = controller_pool(p, states, semantics, obstacles, k, obs_radius, rows)
    q, qg = (states, goals) if rows is None else (states[rows], goals[rows])
    err = q[:, :2] - qg
    vel = q[:, 2:]
    y = torch.cat([z, err, vel], dim=1)                                                  # controller.py:47
    y = F.relu(F.linear(y, p["controller_dec_net.0.weight"], p["controller_dec_net.0.bias"]))
    y = F.relu(F.linear(y, p["controller_dec_net.2.weight"], p["controller_dec_net.2.bias"]))
    y = F.relu(F.linear(y, p["controller_dec_net.4.weight"], p["controller_dec_net.4.bias"]))
    y = F.linear(y, p["controller_dec_net.6.weight"], p["controller_dec_net.6.bias"])
    gains = 2.0 * torch.sigmoid(y) + 0.2                                                 # controller.py:53
    k1, k2, k3, k4 = torch.split(gains, 1, dim=1)
    zero = torch.zeros_like(k1)
    state = torch.cat([err, vel], dim=1)
    ax = torch.sum(state * -torch.cat([k1, zero, k2, zero], dim=1), dim=1, keepdim=True)  # controller.py:56-60
    ay = torch.sum(state * -torch.cat([zero, k3, zero, k4], dim=1), dim=1, keepdim=True)
    return torch.cat([ax, ay], dim=1)


# ----------------------------------------------------------------------------------------
# dynamics (core.py:74-86), Euler step (train.py:70 / core.py:133-134)
# ----------------------------------------------------------------------------------------
def dynamics(states: torch.Tensor, actions: torch.Tensor) -> torch.Tensor:
    return torch.cat([states[:, 2:], actions], dim=1)


def euler_step(states, actions, dt: float = TIME_STEP):
    return states + dynamics(states, actions) * dt


# ----------------------------------------------------------------------------------------
# TTC mask (core.py:187-209, repairs L2)
# ----------------------------------------------------------------------------------------
def ttc_dangerous_mask(states, obstacles=None, k=TOP_K, r=DIST_MIN_THRES, ttc=TIME_TO_COLLISION):
    """bool (N, K_eff, 1)."""
    e, _ = _gather_edges(states, obstacles, k, False)
    x, y, vx, vy, eye = torch.split(e, 1, dim=2)
    x = x + eye
    y = y + eye
    alpha = vx ** 2 + vy ** 2
    beta = 2 * (x * vx + y * vy)
    gamma = x ** 2 + y ** 2 - r ** 2
    dist_dangerous = torch.less(gamma, 0)
    disc = beta ** 2 - 4 * alpha * gamma
    two_pos_roots = torch.logical_and(torch.greater(disc, 0),
                                      torch.logical_and(torch.greater(gamma, 0), torch.less(beta, 0)))
    root_lt_ttc = torch.logical_or(torch.less(-beta - 2 * alpha * ttc, 0),
                                   torch.less((beta + 2 * alpha * ttc) ** 2, disc))
    return torch.logical_or(dist_dangerous, torch.logical_and(two_pos_roots, root_lt_ttc))


def ttc_dangerous_mask_dense_np(s: np.ndarray, r=0.07, ttc=0.1) -> np.ndarray:
    """Dense (N,N,1) evaluation-time check, core.py:212-231 (second cross-check of the
    predicate's algebra; not on the product path)."""
    d = s[:, None, :] - s[None, :, :]
    x, y, vx, vy = [d[:, :, i:i + 1] for i in range(4)]
    ey = np.eye(s.shape[0])[:, :, None]
    x = x + ey
    y = y + ey
    alpha = vx ** 2 + vy ** 2
    beta = 2 * (x * vx + y * vy)
    gamma = x ** 2 + y ** 2 - r ** 2
    disc = beta ** 2 - 4 * alpha * gamma
    two = (disc > 0) & (gamma > 0) & (beta < 0)
    root = (-beta - 2 * alpha * ttc < 0) | ((beta + 2 * alpha * ttc) ** 2 < disc)
    return (gamma < 0) | (two & root)


# ----------------------------------------------------------------------------------------
# losses (core.py:89-184 with repairs L1-L9)
# ----------------------------------------------------------------------------------------
def _hinge_pair(values: torch.Tensor, dang: torch.Tensor, dang_fn, safe_fn, dang_acc_fn, safe_acc_fn):
    v = values.reshape(-1)
    dmask = dang.reshape(-1)
    smask = torch.logical_not(dmask)
    dv = torch.masked_select(v, dmask)
    sv = torch.masked_select(v, smask)
    nd = float(dv.shape[0])
    ns = float(sv.shape[0])
    loss_d = torch.sum(torch.clamp(dang_fn(dv), min=0)) / (1e-5 + nd)
    loss_s = torch.sum(torch.clamp(safe_fn(sv), min=0)) / (1e-5 + ns)
    acc_d = torch.sum(dang_acc_fn(dv).float()) / (1e-5 + nd) if nd > 0 else torch.tensor(-1.0)
    acc_s = torch.sum(safe_acc_fn(sv).float()) / (1e-5 + ns) if ns > 0 else torch.tensor(-1.0)
    return loss_d, loss_s, acc_d, acc_s


def loss_barrier(h, states, obstacles=None, k=TOP_K, r=DIST_MIN_THRES, ttc=TIME_TO_COLLISION):
    """core.py:89-129."""
    dang = ttc_dangerous_mask(states, obstacles, k, r, ttc)
    return _hinge_pair(h, dang,
                       lambda v: v + 1e-3, lambda v: -v + 0.0,
                       lambda v: torch.less_equal(v, 0), lambda v: torch.greater(v, 0))


def loss_derivatives(p_cbf, states, actions, h, semantics=REPAIRED, obstacles=None, k=TOP_K,
                     r=DIST_MIN_THRES, ttc=TIME_TO_COLLISION, dt=TIME_STEP, alpha=ALPHA_CBF,
                     obs_radius=OBS_RADIUS, reuse_indices=False, indices_next=None):
    """core.py:132-171 (repair L8: the CBF is evaluated on states_next).  reuse_indices: h' on S's neighbour sets.
    indices_next (N, K) int64: checker hook - evaluate h' on these neighbour sets of S' instead of a fresh top-k, so
    a test can still compare losses and gradients when a 1-ulp difference in u flipped a near-tie row of S''s graph."""
    states_next = euler_step(states, actions, dt)
    ind = neighbor_indices(states.detach(), k, obstacles).long() if reuse_indices else None
    if indices_next is not None:
        ind = indices_next.long()
    h_next = cbf_forward(p_cbf, states_next, semantics, obstacles, k, obs_radius, ind)
    deriv = h_next - h + dt * alpha * h
    dang = ttc_dangerous_mask(states, obstacles, k, r, ttc)
    out = _hinge_pair(deriv, dang,
                      lambda v: -v + 1e-3, lambda v: -v + 0.0,
                      lambda v: torch.greater_equal(v, 0), lambda v: torch.greater(v, 0))
    return out, h_next


def loss_actions(s, g, a):
    """core.py:174-184."""
    gain = -torch.tensor(np.eye(2, 4) + np.eye(2, 4, k=2) * np.sqrt(3), dtype=torch.float32)
    s_ref = torch.cat([s[:, :2] - g, s[:, 2:]], dim=1)
    a_ref = torch.matmul(s_ref, gain.T.to(s.dtype))
    return torch.mean(torch.abs(torch.sum(torch.square(a), dim=1) - torch.sum(torch.square(a_ref), dim=1)))


def total_loss(l_dang, l_safe, l_dang_d, l_safe_d, l_act):
    """train.py:93,98 (repair T1)."""
    return 10 * (2 * l_dang + l_safe + 2 * l_dang_d + l_safe_d + 0.01 * l_act)


# ----------------------------------------------------------------------------------------
# one full per-env step (the unit the benchmark counts): forward and, optionally, gradients
# ----------------------------------------------------------------------------------------
def full_step(p_cbf, p_ctrl, states, goals, obstacles=None, k=TOP_K, semantics=REPAIRED,
              with_grad=False, r=DIST_MIN_THRES, ttc=TIME_TO_COLLISION, dt=TIME_STEP,
              alpha=ALPHA_CBF, obs_radius=OBS_RADIUS, reuse_indices=False, indices_next=None):
    """graph(S) -> u -> S' -> graph(S') -> h, h' -> TTC mask -> 9 scalars [-> gradients].
    Returns a dict of tensors.  The action fed to the derivative loss keeps its autograd
    link to the controller (train.py:60-70 is not detached)."""
    if with_grad:
        p_cbf = {n: t.detach().clone().requires_grad_(True) for n, t in p_cbf.items()}
        p_ctrl = {n: t.detach().clone().requires_grad_(True) for n, t in p_ctrl.items()}
    ctx = torch.enable_grad() if with_grad else torch.no_grad()
    with ctx:
        u = controller_forward(p_ctrl, states, goals, semantics, obstacles, k, obs_radius)
        h = cbf_forward(p_cbf, states, semantics, obstacles, k, obs_radius)
        lb = loss_barrier(h, states, obstacles, k, r, ttc)
        ld, h_next = loss_derivatives(p_cbf, states, u, h, semantics, obstacles, k, r, ttc, dt, alpha,
                                      obs_radius, reuse_indices, indices_next)
        la = loss_actions(states, goals, u)
        total = total_loss(lb[0], lb[1], ld[0], ld[1], la)
        out = {
            "u": u, "h": h, "h_next": h_next, "s_next": euler_step(states, u, dt),
            "ttc_mask": ttc_dangerous_mask(states, obstacles, k, r, ttc),
            "scalars": torch.stack([lb[0], lb[1], lb[2], lb[3], ld[0], ld[1], ld[2], ld[3], la]),
            "total": total,
        }
        if with_grad:
            total.backward()
            out["grad_cbf"] = torch.cat([p_cbf[n].grad.reshape(-1) if p_cbf[n].grad is not None
                                         else torch.zeros(p_cbf[n].numel()) for n in CBF_KEYS])
            out["grad_ctrl"] = torch.cat([p_ctrl[n].grad.reshape(-1) if p_ctrl[n].grad is not None
                                          else torch.zeros(p_ctrl[n].numel()) for n in CTRL_KEYS])
    return {n: (t.detach() if isinstance(t, torch.Tensor) else t) for n, t in out.items()}


# ----------------------------------------------------------------------------------------
# synthetic swarms - used by tests and bench on both arms
# ----------------------------------------------------------------------------------------
def synthetic_swarm(n_envs: int, n_agents: int, seed: int, vel_std: float = 0.3):
    """positions U[0, sqrt(max(1,N/8)))^2, velocities N(0, vel_std^2), goals p + U(-.5,.5)^2."""
    g = torch.Generator().manual_seed(seed)
    side = math.sqrt(max(1.0, n_agents / 8.0))
    pos = torch.rand(n_envs, n_agents, 2, generator=g) * side
    vel = torch.randn(n_envs, n_agents, 2, generator=g) * vel_std
    goals = pos + (torch.rand(n_envs, n_agents, 2, generator=g) - 0.5)
    return torch.cat([pos, vel], dim=2).contiguous(), goals.contiguous()


def synthetic_obstacles(n_envs: int, n_agents: int, n_circles: int, n_rects: int, seed: int) -> torch.Tensor:
    """Obstacle point sets in the format of core.py:7-42 (12 points per obstacle, float64
    arithmetic, cast to fp32): circles r=0.3, rectangles 0.6 x 0.4."""
    rng = np.random.RandomState(seed)
    side = math.sqrt(max(1.0, n_agents / 8.0))
    envs = []
    for _ in range(n_envs):
        pts = []
        for _ in range(n_circles):
            c = rng.uniform(0, side, size=2)
            th = np.linspace(0, 2 * np.pi, num=12, endpoint=False).reshape(-1, 1)
            pts.append(c + np.concatenate([np.cos(th), np.sin(th)], axis=1) * 0.3)
        for _ in range(n_rects):
            c = rng.uniform(0, side, size=2)
            a, b = 0.6, 0.4
            n1 = int(12 // 2 * a / (a + b)); n2 = 12 // 2 - n1; n3 = n1; n4 = 12 - n1 - n2 - n3
            top = np.stack([np.linspace(-a / 2, a / 2, n1, endpoint=False), np.full(n1, b / 2)], 1)
            right = np.stack([np.full(n2, a / 2), np.linspace(b / 2, -b / 2, n2, endpoint=False)], 1)
            bottom = np.stack([np.linspace(a / 2, -a / 2, n3, endpoint=False), np.full(n3, -b / 2)], 1)
            left = np.stack([np.full(n4, -a / 2), np.linspace(-b / 2, b / 2, n4, endpoint=False)], 1)
            pts.append(np.concatenate([top, right, bottom, left], 0) + c)
        envs.append(np.concatenate(pts, 0) if pts else np.zeros((0, 2)))
    return torch.from_numpy(np.stack(envs, 0).astype(np.float32))


# ----------------------------------------------------------------------------------------
# scenario generation (core.py:45-71) on a counter-based stream - the sequential restatement of what
# csrc/scenario.cu resolves in parallel (same stream, same fp32 decisions, identical bits)
# ----------------------------------------------------------------------------------------
_M64 = (1 << 64) - 1


def _mix64(z: int) -> int:
    z &= _M64
    z ^= z >> 30
    z = (z * 0xBF58476D1CE4E5B9) & _M64
    z ^= z >> 27
    z = (z * 0x94D049BB133111EB) & _M64
    z ^= z >> 31
    return z


def stream_uniform2(seed: int, env: int, stream: int, index: int):
    """Two fp32 uniforms in [0,1) with 24 random bits each (scenario.cu: rng2)."""
    h = _mix64((seed & _M64) ^ _mix64(env + 1))
    h = _mix64(h ^ _mix64(stream + 0x1234567))
    h = _mix64(h + index * 0x9E3779B97F4A7C15)
    scale = np.float32(2.0 ** -24)
    return np.float32(h >> 40) * scale, np.float32((h >> 8) & 0xFFFFFF) * scale


def _dist32(ax, ay, bx, by):
    dx = np.float32(ax) - np.float32(bx)
    dy = np.float32(ay) - np.float32(by)
    return np.sqrt(dx * dx + dy * dy)


def generate_data_stream(seed: int, env: int, n_agents: int, dist_min_thres: float = DIST_MIN_THRES,
                         n_candidates: Optional[int] = None):
    """core.py:45-71 step by step (dart throwing with the unplaced rows at the origin), drawing from the
    counter-based stream: -> (states (N,4) float32, goals (N,2) float32), or None if the candidates run out."""
    thres = np.float32(dist_min_thres)
    side = np.float32(math.sqrt(max(1.0, n_agents / 8.0)))
    T = n_candidates if n_candidates else 2 * n_agents + 256
    pos = np.zeros((n_agents, 2), dtype=np.float32)
    i = 0
    for t in range(T):
        if i == n_agents:
            break
        u0, u1 = stream_uniform2(seed, env, 0, t)
        c = np.array([u0 * side, u1 * side], dtype=np.float32)
        if _dist32(c[0], c[1], 0, 0) <= thres:                        # rows i.. are still (0,0): core.py:47,54
            continue
        if i and np.min(_dist32(pos[:i, 0], pos[:i, 1], c[0], c[1])) <= thres:
            continue
        pos[i] = c
        i += 1
    if i < n_agents:
        return None
    goals = np.zeros((n_agents, 2), dtype=np.float32)
    half = np.float32(0.5)
    for i in range(n_agents):
        a = 0
        while True:
            u0, u1 = stream_uniform2(seed, env, 1, (i << 20) | a)
            g = np.array([(u0 - half) + pos[i, 0], (u1 - half) + pos[i, 1]], dtype=np.float32)   # core.py:59
            ok = _dist32(g[0], g[1], 0, 0) > thres
            if ok and i:
                ok = np.min(_dist32(goals[:i, 0], goals[:i, 1], g[0], g[1])) > thres
            if ok:
                break
            a += 1
        goals[i] = g
    return np.concatenate([pos, np.zeros_like(pos)], axis=1), goals
