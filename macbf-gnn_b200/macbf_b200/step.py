"""One fused training step of the safety-certificate path on a batch of independent envs.

forward  : macbf_step_forward (graph(S) -> u, S' -> graph(S') -> h, h' -> TTC mask -> scalars)
objective: mean over envs of the per-env reference loss
           10 * (2 l_dang + l_safe + 2 l_dang' + l_safe' + 0.01 l_act)        (train.py:93,98)
backward : loss seeds -> CBF backward on (S', graph') and on (S, graph) -> controller backward with
           d/du = d l_act/du + dt * d/dS'[v]  (S' = S + [v,u] dt couples the CBF and controller
           gradients, core.py:133-137)
sync     : all-reduce (sum) of the flat 51,141-float gradient buffer across ranks, after which every
           rank holds the gradient of the global env-mean.  The CBF part [0,17089) is
           finished before the controller backward starts, so its all-reduce is issued there and
           overlaps it; the controller part follows at the end.
All arithmetic runs in libmacbf_b200.so; torch provides buffers, the stream and NCCL.
"""
from typing import Optional

import torch

from . import lib as _lib
from . import ops

# d(objective)/d(scalars) for one env, before the 1/B_global env mean (train.py:93: weights 2,1,2,1,0.01; x10)
_LOSS_WEIGHTS = {0: 20.0, 1: 10.0, 4: 20.0, 5: 10.0, 8: 0.1}


class TrainStep:
    # record_forward=None: the forward pass writes the CBF tile records (one CBF forward per pass less) whenever the two
    # record buffers fit this budget; False: recompute in the backward (macbf_cbf_backward), no extra memory
    RECORD_BUDGET_BYTES = 48 << 30      # of 180 GB HBM; BASELINE C4's per-GPU share (32 envs x 16,384 agents) needs 18.6 GB

    def __init__(self, B: int, N: int, M: int, top_k: int, device, world_size: int = 1, process_group=None,
                 record_forward: Optional[bool] = None):
        self.B, self.N, self.M, self.top_k = B, N, M, top_k
        self.world_size = world_size
        self.group = process_group
        self.device = torch.device(device)
        L = _lib.load()
        K = ops.neighbor_count(N, M, top_k)
        if record_forward is None:
            record_forward = B * N * K < 2 ** 31 and 2 * L.macbf_cbf_record_bytes(B, N, K) <= self.RECORD_BUDGET_BYTES
        self.record_forward = bool(record_forward)
        self.buf = ops.StepBuffers(B, N, M, top_k, device, record=self.record_forward)
        f32 = torch.float32
        dev = device
        # one flat gradient buffer [cbf | ctrl] so a single collective moves everything
        self.grad = torch.zeros((ops.CBF_PARAMS + ops.CTRL_PARAMS,), dtype=f32, device=dev)
        self.grad_cbf = self.grad[:ops.CBF_PARAMS]
        self.grad_ctrl = self.grad[ops.CBF_PARAMS:]
        self.gh = torch.empty((B, N, K), dtype=f32, device=dev)
        self.ghn = torch.empty((B, N, K), dtype=f32, device=dev)
        self.gu = torch.empty((B, N, 2), dtype=f32, device=dev)
        self.gs_next = torch.zeros((B, N, 4), dtype=f32, device=dev)
        gsc = torch.zeros((B, ops.NUM_SCALARS), dtype=f32)
        for c, w in _LOSS_WEIGHTS.items():
            gsc[:, c] = w / (B * world_size)
        self.grad_scalars = gsc.to(dev)
        nb_cbf = L.macbf_cbf_backward_record_scratch_bytes() if self.record_forward else L.macbf_cbf_backward_scratch_bytes(B, N, K)
        nb = max(nb_cbf, L.macbf_controller_backward_scratch_bytes(B, N, K), 256)
        self.scratch = torch.empty((nb,), dtype=torch.uint8, device=dev)
        self.overlap_allreduce = True
        # optional probes (bench.py): recorded on the current stream after the last backward kernel / after the sync
        self.ev_backward_done = None
        self.ev_sync_done = None

    def forward_backward(self, states, goals, cbf_params, ctrl_params, obstacles: Optional[torch.Tensor] = None,
                         all_reduce: bool = True, reuse_indices: bool = False):
        """Returns (scalars [B,12], grad [51141] = [grad_cbf | grad_ctrl] of the env-mean objective).
        reuse_indices: h' on S's neighbour sets (ops.step_forward); the backward needs no change, idx_next = idx."""
        with torch.cuda.device(states.device):     # launch on the device (and stream) that owns the buffers
            return self._forward_backward(states, goals, cbf_params, ctrl_params, obstacles, all_reduce, reuse_indices)

    def _forward_backward(self, states, goals, cbf_params, ctrl_params, obstacles, all_reduce, reuse_indices):
        L = _lib.load()
        b = self.buf
        B, N, M, K = self.B, self.N, self.M, b.K
        ops.step_forward(states, goals, cbf_params, ctrl_params, b, obstacles, reuse_indices=reuse_indices,
                         save_winners=True,   # the controller backward below reuses pool and the max-pool winners
                         record=self.record_forward)
        st = torch.cuda.current_stream().cuda_stream
        p = ops._p
        obs = obstacles if (obstacles is not None and M > 0) else None
        _lib.check(L.macbf_loss_backward(p(states), p(goals), p(b.h), p(b.h_next), p(b.u), p(b.ttc_mask), p(b.scalars),
                                         p(self.grad_scalars), B, N, K, ops.TIME_STEP, ops.ALPHA_CBF, p(self.gh),
                                         p(self.ghn), p(self.gu), None, st), "macbf_loss_backward")
        self.grad.zero_()
        self.gs_next.zero_()
        if self.record_forward:
            # the forward wrote the tile records: the backward starts at the delta chain (h' first: parameter gradient +
            # d/dS'; then h: S is the data of this step, only the parameter gradient is needed)
            _lib.check(L.macbf_cbf_backward_record(p(b.idx_next), p(cbf_params), p(self.ghn), B, N, K, p(b.record_next),
                                                   b.record_bytes, p(self.grad_cbf), p(self.gs_next), p(self.scratch),
                                                   self.scratch.numel(), st), "macbf_cbf_backward_record(S')")
            _lib.check(L.macbf_cbf_backward_record(p(b.idx), p(cbf_params), p(self.gh), B, N, K, p(b.record),
                                                   b.record_bytes, p(self.grad_cbf), None, p(self.scratch),
                                                   self.scratch.numel(), st), "macbf_cbf_backward_record(S)")
        else:
            self._cbf_backward_recompute(L, p, b, states, obs, cbf_params, B, N, M, K, st)
        cbf_work = None
        if all_reduce and self.world_size > 1 and self.overlap_allreduce:
            import torch.distributed as dist
            cbf_work = dist.all_reduce(self.grad_cbf, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return self._controller_backward_and_sync(L, p, b, states, goals, obs, ctrl_params, B, N, M, K, st, all_reduce, cbf_work)

    def _cbf_backward_recompute(self, L, p, b, states, obs, cbf_params, B, N, M, K, st):
        # h' = CBF(S') on its own graph: parameter gradient + d/dS'
        _lib.check(L.macbf_cbf_backward(p(b.s_next), p(obs), p(b.idx_next), p(cbf_params), p(self.ghn), B, N, M, K,
                                        ops.DIST_MIN_THRES, ops.OBS_RADIUS, p(self.grad_cbf), p(self.gs_next),
                                        p(self.scratch), self.scratch.numel(), st), "macbf_cbf_backward(S')")
        # h = CBF(S): S is the data of this step, only the parameter gradient is needed
        _lib.check(L.macbf_cbf_backward(p(states), p(obs), p(b.idx), p(cbf_params), p(self.gh), B, N, M, K,
                                        ops.DIST_MIN_THRES, ops.OBS_RADIUS, p(self.grad_cbf), None,
                                        p(self.scratch), self.scratch.numel(), st), "macbf_cbf_backward(S)")

    def _controller_backward_and_sync(self, L, p, b, states, goals, obs, ctrl_params, B, N, M, K, st, all_reduce, cbf_work):
        # (The CBF half of the flat gradient was final before this call: its all-reduce was started there, on NCCL's own
        # stream, so it runs under the controller backward; the controller half follows after it - SURVEY 8e's one
        # collective, split in two messages so that only the second one is exposed.)
        # controller: d/du from the action loss, plus dt * d/dS'[v] through the Euler step
        ws = b.workspace.data_ptr()
        _lib.check(L.macbf_controller_backward_saved(p(states), p(goals), p(obs), p(b.idx), p(ctrl_params), p(self.gu),
                                                     p(self.gs_next), B, N, M, K, ops.OBS_RADIUS, ops.TIME_STEP, ws,
                                                     ws + b.winners_offset, p(self.grad_ctrl), None, None,
                                                     p(self.scratch), self.scratch.numel(), st),
                   "macbf_controller_backward_saved")
        if self.ev_backward_done is not None:
            self.ev_backward_done.record()
        if all_reduce:
            if cbf_work is not None:
                import torch.distributed as dist
                ctrl_work = dist.all_reduce(self.grad_ctrl, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
                cbf_work.wait()          # stream-level waits: the host does not block
                ctrl_work.wait()
            else:
                sync_gradients(self.grad, self.world_size, self.group)
        if self.ev_sync_done is not None:
            self.ev_sync_done.record()
        return b.scalars, self.grad


def sync_gradients(flat_grad: torch.Tensor, world_size: int, group=None) -> torch.Tensor:
    """The only collective of the path: all-reduce (sum) of the flat [cbf | ctrl] gradient buffer.
    Every rank seeded its backward with 1/B_global, so the sum is the gradient of the global
    env-mean objective.  NCCL on GPUs; the same call runs over gloo in the CPU tests."""
    if world_size > 1:
        import torch.distributed as dist
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
    return flat_grad


def shard_envs(n_envs: int, world_size: int, rank: int):
    """Contiguous env shard [lo, hi) of rank `rank` (envs are independent)."""
    base, rem = divmod(n_envs, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def env_mean_grad_scalars(n_local: int, n_global: int, device=None) -> torch.Tensor:
    """d(objective)/d(scalars) [n_local,12] for the global env-mean objective."""
    g = torch.zeros((n_local, ops.NUM_SCALARS), dtype=torch.float32, device=device)
    for c, w in _LOSS_WEIGHTS.items():
        g[:, c] = w / n_global
    return g
