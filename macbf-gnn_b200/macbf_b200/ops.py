"""Tensor-level wrappers over the C ABI (include/macbf_b200.h).

torch is used for device memory, streams and shapes only; all arithmetic happens in
libmacbf_b200.so.  Every function takes CUDA fp32 tensors with a leading env-batch dim
(states [B,N,4], goals [B,N,2], obstacles [B,M,2] or None) and launches on the current stream.
"""
import functools
from typing import Optional, Tuple

import torch

from . import lib as _lib

LITERAL = 0
REPAIRED = 1
CBF_PARAMS = 17089
CTRL_PARAMS = 34052
NUM_SCALARS = 12

# defaults = the reference's config.py
DIST_MIN_THRES = 0.07
OBS_RADIUS = 1.0
TOP_K = 12
TIME_TO_COLLISION = 2.0
TIME_STEP = 0.1
ALPHA_CBF = 1.0


def _need_cuda(*tensors):
    """All tensor arguments must be CUDA tensors of one device with a supported dtype."""
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise _lib.MacbfError("macbf_b200 has no CPU path: tensors must live on a CUDA device")
        if t.dtype not in (torch.float32, torch.int32, torch.uint8):
            raise _lib.MacbfError(f"unsupported dtype {t.dtype}")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise _lib.MacbfError(f"tensor arguments live on different devices ({dev} and {t.device})")


def _on_device_of(fn):
    """Run `fn` with the device of its first tensor argument current: the kernels launch on that device's current
    stream (not on whichever device happens to be current in the caller), and library state that is per device
    (macbf_loss_forward_noalloc's scratch address) resolves to the device that owns the buffers."""
    @functools.wraps(fn)
    def wrapped(*args, **kwargs):
        for a in list(args) + list(kwargs.values()):
            if isinstance(a, torch.Tensor) and a.is_cuda:
                with torch.cuda.device(a.device):
                    return fn(*args, **kwargs)
        return fn(*args, **kwargs)
    return wrapped


def _c(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    return None if t is None else (t if t.is_contiguous() else t.contiguous())


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _dims(states, obstacles):
    B, N, four = states.shape
    assert four == 4, "states must be [B,N,4]"
    M = 0
    if obstacles is not None:
        assert obstacles.dim() == 3 and obstacles.shape[0] == B and obstacles.shape[2] == 2, "obstacles must be [B,M,2]"
        M = obstacles.shape[1]
        if M == 0:
            obstacles = None
    return B, N, M, obstacles


def neighbor_count(N: int, M: int, top_k: int) -> int:
    return min(top_k, N + M)


@_on_device_of
def neighbor_topk(states: torch.Tensor, top_k: int = TOP_K, obstacles: Optional[torch.Tensor] = None,
                  brute_force: bool = False) -> torch.Tensor:
    """core.py:234-250 -> idx [B,N,K_eff] int32 (uniform-grid search when it applies, else brute force)."""
    _need_cuda(states, obstacles)
    states, obstacles = _c(states), _c(obstacles)
    B, N, M, obstacles = _dims(states, obstacles)
    K = neighbor_count(N, M, top_k)
    idx = torch.empty((B, N, K), dtype=torch.int32, device=states.device)
    L = _lib.load()
    nbytes = L.macbf_neighbor_workspace_bytes(B, N, M, top_k)
    if nbytes == 0 or brute_force:
        _lib.check(L.macbf_neighbor_topk(_p(states), _p(obstacles), B, N, M, top_k, _p(idx), _stream()),
                   "macbf_neighbor_topk")
    else:
        ws = torch.empty((nbytes,), dtype=torch.uint8, device=states.device)
        _lib.check(L.macbf_neighbor_topk_ws(_p(states), _p(obstacles), B, N, M, top_k, _p(idx), _p(ws), nbytes,
                                            _stream()), "macbf_neighbor_topk_ws")
    return idx


@_on_device_of
def neighbor_topk_part(states, top_k, part, nparts, obstacles=None, out=None):
    """Rows of the neighbour graph owned by part `part` of `nparts` (macbf_neighbor_topk_part); the other rows of
    `out` (default: zeros) are left as they are.  Summing the results of all parts gives neighbor_topk."""
    _need_cuda(states, obstacles)
    states, obstacles = _c(states), _c(obstacles)
    B, N, M, obstacles = _dims(states, obstacles)
    K = neighbor_count(N, M, top_k)
    if out is None:
        out = torch.zeros((B, N, K), dtype=torch.int32, device=states.device)
    L = _lib.load()
    nbytes = L.macbf_neighbor_workspace_bytes(B, N, M, top_k)
    ws = torch.empty((max(nbytes, 16),), dtype=torch.uint8, device=states.device)
    _lib.check(L.macbf_neighbor_topk_part(_p(states), _p(obstacles), B, N, M, top_k, part, nparts, _p(out), _p(ws), nbytes,
                                          _stream()), "macbf_neighbor_topk_part")
    return out


def neighbor_topk_sharded(states, top_k, obstacles=None, group=None):
    """One large env row-sharded over the ranks of `group` (positions replicated): every rank searches its share of
    the queries, one all-reduce(sum) of the int32 index array assembles the full graph on every rank."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank(group) if world > 1 else 0
    idx = neighbor_topk_part(states, top_k, rank, world, obstacles)
    if world > 1:
        dist.all_reduce(idx, op=dist.ReduceOp.SUM, group=group)
    return idx


@_on_device_of
def dense_topk(x: torch.Tensor, k: int) -> torch.Tensor:
    """core.remove_distant_agents' selection on a materialised pair tensor x [N,C,F] -> idx [N,k] int32."""
    _need_cuda(x)
    x = _c(x)
    n, c, f = x.shape
    idx = torch.empty((n, k), dtype=torch.int32, device=x.device)
    L = _lib.load()
    _lib.check(L.macbf_dense_topk(_p(x), n, c, f, k, _p(idx), _stream()), "macbf_dense_topk")
    return idx


@_on_device_of
def cbf_forward(states, idx, cbf_params, obstacles=None, semantics=REPAIRED, dist_min_thres=DIST_MIN_THRES,
                obs_radius=OBS_RADIUS) -> torch.Tensor:
    """cbf.py:21-45 -> h [B,N,K]."""
    _need_cuda(states, idx, cbf_params, obstacles)
    states, idx, cbf_params, obstacles = _c(states), _c(idx), _c(cbf_params), _c(obstacles)
    B, N, M, obstacles = _dims(states, obstacles)
    K = idx.shape[2]
    assert cbf_params.numel() == CBF_PARAMS and idx.dtype == torch.int32
    h = torch.empty((B, N, K), dtype=torch.float32, device=states.device)
    L = _lib.load()
    _lib.check(L.macbf_cbf_forward(_p(states), _p(obstacles), _p(idx), _p(cbf_params), B, N, M, K, semantics,
                                   dist_min_thres, obs_radius, _p(h), _stream()), "macbf_cbf_forward")
    return h


@_on_device_of
def controller_forward(states, goals, idx, ctrl_params, obstacles=None, semantics=REPAIRED, obs_radius=OBS_RADIUS,
                       dt=TIME_STEP, want_next=True) -> Tuple[torch.Tensor, Optional[torch.Tensor], torch.Tensor]:
    """controller.py:31-63 + train.py:70 -> (u [B,N,2], s_next [B,N,4] | None, pool [B,N,128])."""
    _need_cuda(states, goals, idx, ctrl_params, obstacles)
    states, goals, idx, ctrl_params, obstacles = _c(states), _c(goals), _c(idx), _c(ctrl_params), _c(obstacles)
    B, N, M, obstacles = _dims(states, obstacles)
    K = idx.shape[2]
    assert ctrl_params.numel() == CTRL_PARAMS and idx.dtype == torch.int32
    dev = states.device
    pool = torch.empty((B, N, 128), dtype=torch.float32, device=dev)
    u = torch.empty((B, N, 2), dtype=torch.float32, device=dev)
    s_next = torch.empty((B, N, 4), dtype=torch.float32, device=dev) if want_next else None
    L = _lib.load()
    _lib.check(L.macbf_controller_forward(_p(states), _p(goals), _p(obstacles), _p(idx), _p(ctrl_params), B, N, M, K,
                                          semantics, obs_radius, dt, _p(pool), _p(u), _p(s_next), _stream()),
               "macbf_controller_forward")
    return u, s_next, pool


@_on_device_of
def ttc_mask(states, idx, obstacles=None, dist_min_thres=DIST_MIN_THRES, ttc=TIME_TO_COLLISION) -> torch.Tensor:
    """core.py:187-209 -> uint8 [B,N,K]."""
    _need_cuda(states, idx, obstacles)
    states, idx, obstacles = _c(states), _c(idx), _c(obstacles)
    B, N, M, obstacles = _dims(states, obstacles)
    K = idx.shape[2]
    mask = torch.empty((B, N, K), dtype=torch.uint8, device=states.device)
    L = _lib.load()
    _lib.check(L.macbf_ttc_mask(_p(states), _p(obstacles), _p(idx), B, N, M, K, dist_min_thres, ttc, _p(mask),
                                _stream()), "macbf_ttc_mask")
    return mask


@_on_device_of
def loss_forward(states, idx, h, h_next=None, u=None, goals=None, obstacles=None, dist_min_thres=DIST_MIN_THRES,
                 ttc=TIME_TO_COLLISION, dt=TIME_STEP, alpha=ALPHA_CBF):
    """core.py:89-184 -> (scalars [B,12], ttc_mask uint8 [B,N,K]).  Scalar layout: macbf_b200.h."""
    _need_cuda(states, idx, h, h_next, u, goals, obstacles)
    states, idx, h, h_next, u, goals, obstacles = map(_c, (states, idx, h, h_next, u, goals, obstacles))
    B, N, M, obstacles = _dims(states, obstacles)
    K = idx.shape[2]
    dev = states.device
    L = _lib.load()
    nbytes = L.macbf_loss_scratch_bytes(B, N, K)
    scratch = torch.empty((max(nbytes, 256),), dtype=torch.uint8, device=dev)
    mask = torch.empty((B, N, K), dtype=torch.uint8, device=dev)
    scalars = torch.empty((B, NUM_SCALARS), dtype=torch.float32, device=dev)
    _lib.check(L.macbf_loss_forward(_p(states), _p(goals), _p(obstacles), _p(idx), _p(h), _p(h_next), _p(u), B, N, M,
                                    K, dist_min_thres, ttc, dt, alpha, _p(mask), _p(scalars), _p(scratch),
                                    scratch.numel(), _stream()), "macbf_loss_forward")
    return scalars, mask


@_on_device_of
def loss_backward(states, goals, h, h_next, u, mask, scalars, grad_scalars, dt=TIME_STEP, alpha=ALPHA_CBF,
                  want_h=True, want_h_next=True, want_u=True, want_states=False):
    """Seeds dL/dh, dL/dh_next [B,N,K] and dL/du [B,N,2] from dL/dscalars [B,12]."""
    _need_cuda(states, goals, h, h_next, u, mask, scalars, grad_scalars)
    states, goals, h, h_next, u, mask, scalars, grad_scalars = map(
        _c, (states, goals, h, h_next, u, mask, scalars, grad_scalars))
    B, N, K = h.shape
    dev = h.device
    gh = torch.empty_like(h) if want_h else None
    ghn = torch.empty_like(h) if (want_h_next and h_next is not None) else None
    gu = torch.empty((B, N, 2), dtype=torch.float32, device=dev) if (want_u and u is not None) else None
    gst = torch.empty((B, N, 4), dtype=torch.float32, device=dev) if (want_states and u is not None) else None
    L = _lib.load()
    _lib.check(L.macbf_loss_backward(_p(states), _p(goals), _p(h), _p(h_next), _p(u), _p(mask), _p(scalars),
                                     _p(grad_scalars), B, N, K, dt, alpha, _p(gh), _p(ghn), _p(gu), _p(gst),
                                     _stream()), "macbf_loss_backward")
    return gh, ghn, gu, gst


@_on_device_of
def cbf_backward(states, idx, cbf_params, grad_h, obstacles=None, dist_min_thres=DIST_MIN_THRES,
                 obs_radius=OBS_RADIUS, grad_params=None, want_grad_states=False):
    """Autograd of cbf.py:21-45 -> (grad_params [17089] (accumulated if given), grad_states [B,N,4] | None)."""
    _need_cuda(states, idx, cbf_params, grad_h, obstacles, grad_params)
    states, idx, cbf_params, grad_h, obstacles = map(_c, (states, idx, cbf_params, grad_h, obstacles))
    B, N, M, obstacles = _dims(states, obstacles)
    K = idx.shape[2]
    dev = states.device
    if grad_params is None:
        grad_params = torch.zeros((CBF_PARAMS,), dtype=torch.float32, device=dev)
    gs = torch.zeros((B, N, 4), dtype=torch.float32, device=dev) if want_grad_states else None
    L = _lib.load()
    nbytes = L.macbf_cbf_backward_scratch_bytes(B, N, K)
    scratch = torch.empty((max(nbytes, 256),), dtype=torch.uint8, device=dev)
    _lib.check(L.macbf_cbf_backward(_p(states), _p(obstacles), _p(idx), _p(cbf_params), _p(grad_h), B, N, M, K,
                                    dist_min_thres, obs_radius, _p(grad_params), _p(gs), _p(scratch), scratch.numel(),
                                    _stream()), "macbf_cbf_backward")
    return grad_params, gs


def cbf_record_bytes(B: int, N: int, K: int) -> int:
    """Bytes of the record buffer of cbf_forward_record (173 KB per 128 edge rows)."""
    return int(_lib.load().macbf_cbf_record_bytes(B, N, K))


@_on_device_of
def cbf_forward_record(states, idx, cbf_params, obstacles=None, dist_min_thres=DIST_MIN_THRES, obs_radius=OBS_RADIUS,
                       record=None):
    """h = CBF(states) as cbf_forward (REPAIRED), also writing the tile records cbf_backward_record starts from
    (a training step then runs the CBF forward once instead of twice).  -> (h [B,N,K], record uint8 buffer)."""
    _need_cuda(states, idx, cbf_params, obstacles, record)
    states, idx, cbf_params, obstacles = map(_c, (states, idx, cbf_params, obstacles))
    B, N, M, obstacles = _dims(states, obstacles)
    K = idx.shape[2]
    L = _lib.load()
    nbytes = L.macbf_cbf_record_bytes(B, N, K)
    if record is None:
        record = torch.empty((nbytes,), dtype=torch.uint8, device=states.device)
    h = torch.empty((B, N, K), dtype=torch.float32, device=states.device)
    _lib.check(L.macbf_cbf_forward_record(_p(states), _p(obstacles), _p(idx), _p(cbf_params), B, N, M, K, dist_min_thres,
                                          obs_radius, _p(h), _p(record), record.numel(), _stream()), "macbf_cbf_forward_record")
    return h, record


@_on_device_of
def cbf_backward_record(idx, cbf_params, grad_h, record, grad_params=None, want_grad_states=False):
    """Backward of cbf_forward_record's pass from its record (reusable) -> (grad_params [17089], grad_states | None)."""
    _need_cuda(idx, cbf_params, grad_h, record, grad_params)
    idx, cbf_params, grad_h = map(_c, (idx, cbf_params, grad_h))
    B, N, K = idx.shape
    dev = idx.device
    if grad_params is None:
        grad_params = torch.zeros((CBF_PARAMS,), dtype=torch.float32, device=dev)
    gs = torch.zeros((B, N, 4), dtype=torch.float32, device=dev) if want_grad_states else None
    L = _lib.load()
    nbytes = L.macbf_cbf_backward_record_scratch_bytes()
    scratch = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    _lib.check(L.macbf_cbf_backward_record(_p(idx), _p(cbf_params), _p(grad_h), B, N, K, _p(record), record.numel(),
                                           _p(grad_params), _p(gs), _p(scratch), nbytes, _stream()),
               "macbf_cbf_backward_record")
    return grad_params, gs


@_on_device_of
def controller_backward(states, goals, idx, ctrl_params, grad_u, grad_s_next=None, obstacles=None,
                        obs_radius=OBS_RADIUS, dt=TIME_STEP, grad_params=None, want_grad_states=False,
                        want_grad_goals=False):
    """Autograd of controller.py:31-63 (+ Euler step) -> (grad_params [34052], grad_states | None, grad_goals | None)."""
    _need_cuda(states, goals, idx, ctrl_params, grad_u, grad_s_next, obstacles, grad_params)
    states, goals, idx, ctrl_params, grad_u, grad_s_next, obstacles = map(
        _c, (states, goals, idx, ctrl_params, grad_u, grad_s_next, obstacles))
    B, N, M, obstacles = _dims(states, obstacles)
    K = idx.shape[2]
    dev = states.device
    if grad_params is None:
        grad_params = torch.zeros((CTRL_PARAMS,), dtype=torch.float32, device=dev)
    gs = torch.zeros((B, N, 4), dtype=torch.float32, device=dev) if want_grad_states else None
    gg = torch.zeros((B, N, 2), dtype=torch.float32, device=dev) if want_grad_goals else None
    L = _lib.load()
    nbytes = L.macbf_controller_backward_scratch_bytes(B, N, K)
    scratch = torch.empty((max(nbytes, 256),), dtype=torch.uint8, device=dev)
    _lib.check(L.macbf_controller_backward(_p(states), _p(goals), _p(obstacles), _p(idx), _p(ctrl_params), _p(grad_u),
                                           _p(grad_s_next), B, N, M, K, obs_radius, dt, _p(grad_params), _p(gs),
                                           _p(gg), _p(scratch), scratch.numel(), _stream()),
               "macbf_controller_backward")
    return grad_params, gs, gg


STEP_REUSE_INDICES = 1   # include/macbf_b200.h: MACBF_STEP_REUSE_INDICES
STEP_SAVE_WINNERS = 2    # MACBF_STEP_SAVE_WINNERS


class StepBuffers:
    """Pre-allocated outputs + workspace for macbf_step_forward: the timed loop allocates nothing."""

    def __init__(self, B: int, N: int, M: int, top_k: int, device, record: bool = False):
        L = _lib.load()
        self.B, self.N, self.M, self.top_k = B, N, M, top_k
        self.K = K = neighbor_count(N, M, top_k)
        # record=True: tile records of the two CBF passes (training: step_forward(..., record=True), then
        # macbf_cbf_backward_record instead of macbf_cbf_backward)
        self.record_bytes = int(L.macbf_cbf_record_bytes(B, N, K)) if record else 0
        self.record = torch.empty((self.record_bytes,), dtype=torch.uint8, device=device) if record else None
        self.record_next = torch.empty((self.record_bytes,), dtype=torch.uint8, device=device) if record else None
        f32, dev = torch.float32, device
        self.idx = torch.empty((B, N, K), dtype=torch.int32, device=dev)
        self.idx_next = torch.empty((B, N, K), dtype=torch.int32, device=dev)
        # `result` = [scalars | u] in ONE block, so a host that wants the losses and the actions of a step reads
        # them back with a single device-to-host copy
        self.result = torch.empty((B * NUM_SCALARS + B * N * 2,), dtype=f32, device=dev)
        self.scalars = self.result[:B * NUM_SCALARS].view(B, NUM_SCALARS)
        self.u = self.result[B * NUM_SCALARS:].view(B, N, 2)
        self.s_next = torch.empty((B, N, 4), dtype=f32, device=dev)
        self.h = torch.empty((B, N, K), dtype=f32, device=dev)
        self.h_next = torch.empty((B, N, K), dtype=f32, device=dev)
        self.ttc_mask = torch.empty((B, N, K), dtype=torch.uint8, device=dev)
        self.workspace = torch.empty((L.macbf_step_workspace_bytes(B, N, M, top_k),), dtype=torch.uint8, device=dev)
        self.winners_offset = L.macbf_step_workspace_winners_offset(B, N, M, top_k)   # valid after save_winners=True


@_on_device_of
def step_forward(states, goals, cbf_params, ctrl_params, buf: StepBuffers, obstacles=None, semantics=REPAIRED,
                 dist_min_thres=DIST_MIN_THRES, obs_radius=OBS_RADIUS, ttc=TIME_TO_COLLISION, dt=TIME_STEP,
                 alpha=ALPHA_CBF, reuse_indices=False, save_winners=False, record=False):
    """The whole per-step path (graph -> u, S' -> graph' -> h, h' -> TTC mask -> scalars) in one C call.
    reuse_indices=True evaluates h' on S's neighbour sets (MACBF_STEP_REUSE_INDICES; not the reference's default).
    record=True (buf made with record=True): the two CBF passes write their tile records (macbf_step_forward_train)."""
    _need_cuda(states, goals, cbf_params, ctrl_params, obstacles)
    B, N, M, obstacles = _dims(states, obstacles)
    assert (B, N, M) == (buf.B, buf.N, buf.M), "StepBuffers were sized for another problem"
    assert states.is_contiguous() and goals.is_contiguous() and cbf_params.is_contiguous() and ctrl_params.is_contiguous()
    L = _lib.load()
    flags = (STEP_REUSE_INDICES if reuse_indices else 0) | (STEP_SAVE_WINNERS if save_winners else 0)
    if record:
        assert buf.record is not None, "StepBuffers were made without record=True"
        _lib.check(L.macbf_step_forward_train(_p(states), _p(goals), _p(obstacles), _p(cbf_params), _p(ctrl_params), B, N, M,
                                              buf.top_k, semantics, dist_min_thres, obs_radius, ttc, dt, alpha, flags,
                                              _p(buf.idx), _p(buf.idx_next), _p(buf.u), _p(buf.s_next), _p(buf.h),
                                              _p(buf.h_next), _p(buf.ttc_mask), _p(buf.scalars), _p(buf.workspace),
                                              buf.workspace.numel(), _p(buf.record), _p(buf.record_next), buf.record_bytes,
                                              _stream()), "macbf_step_forward_train")
        return buf
    _lib.check(L.macbf_step_forward_ex(_p(states), _p(goals), _p(obstacles), _p(cbf_params), _p(ctrl_params), B, N, M,
                                       buf.top_k, semantics, dist_min_thres, obs_radius, ttc, dt, alpha, flags,
                                       _p(buf.idx), _p(buf.idx_next),
                                       _p(buf.u), _p(buf.s_next), _p(buf.h), _p(buf.h_next), _p(buf.ttc_mask),
                                       _p(buf.scalars), _p(buf.workspace), buf.workspace.numel(), _stream()),
               "macbf_step_forward")
    return buf


def generate_scenario(seed, B, N, dist_min_thres=DIST_MIN_THRES, device="cuda", n_candidates=None):
    """core.py:45-71 on the device -> (states [B,N,4] with zero velocity, goals [B,N,2]); deterministic in
    (seed, env).  Retries with more candidates in the (unlikely) case an env ran short."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise _lib.MacbfError("macbf_b200 has no CPU path: scenarios are generated on a CUDA device")
    L = _lib.load()
    T = int(n_candidates) if n_candidates else 2 * N + 256
    states = torch.empty((B, N, 4), dtype=torch.float32, device=dev)
    goals = torch.empty((B, N, 2), dtype=torch.float32, device=dev)
    accepted = torch.empty((B,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        for _ in range(6):
            nbytes = L.macbf_scenario_workspace_bytes(B, N, T)
            ws = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
            _lib.check(L.macbf_generate_scenario(int(seed) & 0xFFFFFFFFFFFFFFFF, B, N, T, dist_min_thres, _p(states), _p(goals),
                                                 _p(accepted), _p(ws), nbytes, _stream()), "macbf_generate_scenario")
            if int(accepted.min()) >= N:
                return states, goals
            T *= 2
    raise _lib.MacbfError(f"generate_scenario: could not place {N} agents (threshold {dist_min_thres})")


@_on_device_of
def ttc_dense(states, dist_min_check=0.07, ttc_check=0.1, want_mask=True, want_counts=True):
    """core.py:212-231 on the device: (counts [B,N] int32 | None, mask [B,N,N] uint8 | None), bit-exact."""
    _need_cuda(states)
    states = _c(states)
    B, N, _ = states.shape
    dev = states.device
    counts = torch.empty((B, N), dtype=torch.int32, device=dev) if want_counts else None
    mask = torch.empty((B, N, N), dtype=torch.uint8, device=dev) if want_mask else None
    L = _lib.load()
    _lib.check(L.macbf_ttc_dense(_p(states), B, N, dist_min_check ** 2, ttc_check, _p(counts), _p(mask), _stream()),
               "macbf_ttc_dense")
    return counts, mask


def generate_obstacles(specs, num=12, device="cuda"):
    """core.py:7-42 on the device.  specs: sequence of ("circle", (cx, cy), radius) / ("rectangle", (cx, cy), (a, b)).
    -> float64 [len(specs) * num, 2] point set in the reference's order (cast to float32 for `obstacles=`)."""
    dev = torch.device(device)
    if dev.type != "cuda":
        raise _lib.MacbfError("macbf_b200 has no CPU path: obstacle point sets are generated on a CUDA device")
    kinds, centers, sizes = [], [], []
    for kind, center, size in specs:
        if kind == "circle":
            kinds.append(0); sizes.append((float(size), 0.0))
        elif kind == "rectangle":
            kinds.append(1); sizes.append((float(size[0]), float(size[1])))
        else:
            raise ValueError(f"unknown obstacle kind {kind!r}")
        centers.append((float(center[0]), float(center[1])))
    n = len(kinds)
    k = torch.tensor(kinds, dtype=torch.int32, device=dev)
    c = torch.tensor(centers, dtype=torch.float64, device=dev)
    z = torch.tensor(sizes, dtype=torch.float64, device=dev)
    out = torch.empty((n * num, 2), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.load().macbf_generate_obstacles(_p(k), _p(c), _p(z), n, int(num), _p(out), _stream()),
                   "macbf_generate_obstacles")
    return out


@_on_device_of
def rollout_forward(states, goals, ctrl_params, max_steps, obstacles=None, noise=None, top_k=TOP_K, semantics=REPAIRED,
                    obs_radius=OBS_RADIUS, dt=TIME_STEP, dist_stop=0.07):
    """train.py:59-81 in one call, stop test on the device (no per-step host sync).
    -> (traj_states [T+1,B,N,4], traj_actions [T,B,N,2], steps_run [B] int32)."""
    _need_cuda(states, goals, ctrl_params, obstacles, noise)
    states, goals, ctrl_params, obstacles, noise = map(_c, (states, goals, ctrl_params, obstacles, noise))
    B, N, M, obstacles = _dims(states, obstacles)
    if noise is not None and tuple(noise.shape) != (max_steps, B, N, 2):
        raise ValueError(f"noise must be [{max_steps}, {B}, {N}, 2], got {tuple(noise.shape)}")
    dev = states.device
    L = _lib.load()
    ts = torch.zeros((max_steps + 1, B, N, 4), dtype=torch.float32, device=dev)
    ta = torch.zeros((max_steps, B, N, 2), dtype=torch.float32, device=dev)
    steps = torch.empty((B,), dtype=torch.int32, device=dev)
    nbytes = L.macbf_rollout_workspace_bytes(B, N, M, top_k)
    ws = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
    _lib.check(L.macbf_rollout_forward(_p(states), _p(goals), _p(obstacles), _p(ctrl_params), _p(noise), B, N, M, top_k, max_steps,
                                       semantics, obs_radius, dt, dist_stop, _p(ts), _p(ta), _p(steps), _p(ws), nbytes,
                                       _stream()), "macbf_rollout_forward")
    return ts, ta, steps


@_on_device_of
def rollout_backward(traj_states, goals, ctrl_params, grad_traj_actions, grad_traj_states=None, obstacles=None, top_k=TOP_K,
                     obs_radius=OBS_RADIUS, dt=TIME_STEP, grad_params=None, want_grad_states0=False, want_grad_goals=False):
    """Autograd of the rollout (train.py:59-81 is not detached): dL/da_t [T,B,N,2] and dL/ds_t [T,B,N,4] (optional), zero
    beyond an env's steps_run -> (grad_params [34052], dL/ds_0 | None, grad_goals | None).  traj_states: the first T (or
    more) slots of rollout_forward's trajectory."""
    _need_cuda(traj_states, goals, ctrl_params, grad_traj_actions, grad_traj_states, obstacles, grad_params)
    traj_states, goals, ctrl_params, grad_traj_actions, grad_traj_states, obstacles = map(
        _c, (traj_states, goals, ctrl_params, grad_traj_actions, grad_traj_states, obstacles))
    T = grad_traj_actions.shape[0]
    B, N, M, obstacles = _dims(traj_states[0], obstacles)
    if traj_states.shape[0] < T or tuple(grad_traj_actions.shape) != (T, B, N, 2):
        raise ValueError(f"grad_traj_actions must be [T<={traj_states.shape[0]}, {B}, {N}, 2], got {tuple(grad_traj_actions.shape)}")
    if grad_traj_states is not None and tuple(grad_traj_states.shape) != (T, B, N, 4):
        raise ValueError(f"grad_traj_states must be [{T}, {B}, {N}, 4], got {tuple(grad_traj_states.shape)}")
    dev = traj_states.device
    if grad_params is None:
        grad_params = torch.zeros((CTRL_PARAMS,), dtype=torch.float32, device=dev)
    gs0 = torch.empty((B, N, 4), dtype=torch.float32, device=dev) if want_grad_states0 else None
    gg = torch.zeros((B, N, 2), dtype=torch.float32, device=dev) if want_grad_goals else None
    L = _lib.load()
    nbytes = L.macbf_rollout_backward_scratch_bytes(B, N, M, top_k)
    scratch = torch.empty((max(nbytes, 256),), dtype=torch.uint8, device=dev)
    _lib.check(L.macbf_rollout_backward(_p(traj_states), _p(goals), _p(obstacles), _p(ctrl_params), _p(grad_traj_states),
                                        _p(grad_traj_actions), B, N, M, top_k, T, obs_radius, dt, _p(grad_params), _p(gs0),
                                        _p(gg), _p(scratch), scratch.numel(), _stream()), "macbf_rollout_backward")
    return grad_params, gs0, gg


def launch_count() -> int:
    return int(_lib.load().macbf_launch_count())


def timing_begin():
    """Start recording one CUDA event per library launch on the current stream (macbf_timing_begin)."""
    _lib.check(_lib.load().macbf_timing_begin(_stream()), "macbf_timing_begin")


def timing_end():
    """-> [(kernel name, ms since the previous launch mark), ...] in launch order (macbf_timing_end; synchronises)."""
    import ctypes
    n_max, stride = 4096, 48
    names = ctypes.create_string_buffer(n_max * stride)
    ms = (ctypes.c_float * n_max)()
    rc = _lib.load().macbf_timing_end(n_max, names, stride, ctypes.cast(ms, ctypes.c_void_p))
    if rc > 0:
        _lib.check(rc, "macbf_timing_end")
    raw = names.raw
    return [(raw[i * stride:(i + 1) * stride].split(b"\0", 1)[0].decode(), float(ms[i])) for i in range(-rc)]
