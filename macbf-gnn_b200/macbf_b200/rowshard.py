"""One LARGE env row-sharded over several GPUs: the forward step of `ops.step_forward` for B = 1 with
the query rows (agents) split into contiguous ranges, one per rank; positions are replicated.

    rank r:  rows_a   graph(S) rows -> controller -> u, s_next rows
             all-gather of the s_next rows (16 B per agent)
             rows_b   graph(S') rows -> CBF(S), CBF(S') rows -> TTC mask rows, loss partial sums
             all-reduce (sum) of the partial sums (a few KB)
             rows_c   12 scalars (identical on every rank)

Every rank ends with its own rows of idx / idx_next / u / h / h_next / ttc_mask, the complete s_next and the scalars.
All arithmetic is in libmacbf_b200.so (macbf_step_forward_rows_a/_b/_c); torch provides the buffers, the stream and
the two collectives.  Forward only: the training step shards env batches instead (step.py).
"""
from typing import List, Optional, Tuple

import torch

from . import lib as _lib
from . import ops

ROW_SHARD_ALIGN = 1024   # include/macbf_b200.h: MACBF_ROW_SHARD_ALIGN


def row_ranges(N: int, nparts: int) -> List[Tuple[int, int]]:
    """Contiguous agent ranges, one per part, with bounds on multiples of ROW_SHARD_ALIGN (the last one ends at N).
    Parts beyond the number of aligned blocks get empty ranges."""
    blocks = (N + ROW_SHARD_ALIGN - 1) // ROW_SHARD_ALIGN
    out = []
    for p in range(nparts):
        lo = min(N, (blocks * p // nparts) * ROW_SHARD_ALIGN)
        hi = min(N, (blocks * (p + 1) // nparts) * ROW_SHARD_ALIGN)
        out.append((lo, hi))
    return out


class RowShardedStep:
    """Buffers + the three phases for one env of N agents.  `group` None and world_size 1 = the unsharded step."""

    def __init__(self, N: int, M: int, top_k: int, device, group=None):
        import torch.distributed as dist
        self.N, self.M, self.top_k = N, M, top_k
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        self.ranges = row_ranges(N, self.world)
        self.lo, self.hi = self.ranges[self.rank]
        self.buf = ops.StepBuffers(1, N, M, top_k, device)
        L = _lib.load()
        nb = L.macbf_loss_scratch_bytes(1, N, self.buf.K)
        self.loss_partial = torch.zeros((max(nb, 4) // 4,), dtype=torch.float32, device=device)
        self.equal = len({hi - lo for lo, hi in self.ranges}) == 1

    def phase_a(self, states, goals, ctrl_params, obstacles, lo, hi):
        L, b = _lib.load(), self.buf
        _lib.check(L.macbf_step_forward_rows_a(ops._p(states), ops._p(goals), ops._p(obstacles), ops._p(ctrl_params), self.N,
                                               self.M, self.top_k, ops.REPAIRED, ops.OBS_RADIUS, ops.TIME_STEP, lo, hi,
                                               ops._p(b.idx), ops._p(b.u), ops._p(b.s_next), ops._p(b.workspace),
                                               b.workspace.numel(), ops._stream()), "macbf_step_forward_rows_a")

    def phase_b(self, states, goals, cbf_params, obstacles, lo, hi, loss_partial=None):
        L, b = _lib.load(), self.buf
        lp = self.loss_partial if loss_partial is None else loss_partial
        _lib.check(L.macbf_step_forward_rows_b(ops._p(states), ops._p(b.s_next), ops._p(goals), ops._p(obstacles),
                                               ops._p(cbf_params), self.N, self.M, self.top_k, ops.REPAIRED, ops.DIST_MIN_THRES,
                                               ops.OBS_RADIUS, ops.TIME_TO_COLLISION, ops.TIME_STEP, ops.ALPHA_CBF, lo, hi,
                                               ops._p(b.idx), ops._p(b.idx_next), ops._p(b.u), ops._p(b.h), ops._p(b.h_next),
                                               ops._p(b.ttc_mask), ops._p(lp), ops._p(b.workspace), b.workspace.numel(),
                                               ops._stream()), "macbf_step_forward_rows_b")

    def phase_c(self, loss_partial=None):
        L, b = _lib.load(), self.buf
        lp = self.loss_partial if loss_partial is None else loss_partial
        _lib.check(L.macbf_step_forward_rows_c(ops._p(lp), self.N, b.K, ops._p(b.scalars), ops._stream()),
                   "macbf_step_forward_rows_c")

    def forward(self, states, goals, cbf_params, ctrl_params, obstacles: Optional[torch.Tensor] = None):
        """states [1,N,4], goals [1,N,2] replicated on every rank.  Returns the StepBuffers (own rows + scalars)."""
        ops._need_cuda(states, goals, cbf_params, ctrl_params, obstacles)
        assert tuple(states.shape) == (1, self.N, 4) and states.is_contiguous() and goals.is_contiguous()
        obs = obstacles if (obstacles is not None and self.M > 0) else None
        with torch.cuda.device(states.device):
            self.phase_a(states, goals, ctrl_params, obs, self.lo, self.hi)
            if self.world > 1:
                import torch.distributed as dist
                sn = self.buf.s_next.view(self.N, 4)
                if self.equal:
                    dist.all_gather_into_tensor(sn, sn[self.lo:self.hi].clone(), group=self.group)
                else:       # ragged ranges: zero the foreign rows and add
                    sn[:self.lo].zero_()
                    sn[self.hi:].zero_()
                    dist.all_reduce(sn, op=dist.ReduceOp.SUM, group=self.group)
            self.phase_b(states, goals, cbf_params, obs, self.lo, self.hi)
            if self.world > 1:
                import torch.distributed as dist
                dist.all_reduce(self.loss_partial, op=dist.ReduceOp.SUM, group=self.group)
            self.phase_c()
        return self.buf


def forward_parts_on_one_gpu(step: RowShardedStep, nparts: int, states, goals, cbf_params, ctrl_params, obstacles=None):
    """The row-sharded schedule with `nparts` logical ranks executed one after the other on ONE GPU (tests): the
    all-gather is the shared s_next buffer, the all-reduce a sum of the parts' partial-sum arrays."""
    obs = obstacles if (obstacles is not None and step.M > 0) else None
    ranges = row_ranges(step.N, nparts)
    for lo, hi in ranges:
        step.phase_a(states, goals, ctrl_params, obs, lo, hi)
    total = torch.zeros_like(step.loss_partial)
    for lo, hi in ranges:
        part = torch.empty_like(step.loss_partial)
        step.phase_b(states, goals, cbf_params, obs, lo, hi, loss_partial=part)
        total += part
    step.phase_c(loss_partial=total)
    return step.buf
