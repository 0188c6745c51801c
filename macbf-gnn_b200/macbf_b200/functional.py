"""torch.autograd bridges: the reference's call surface (cbf.py / controller.py / core.py) expressed
as autograd Functions whose forward AND backward run in libmacbf_b200.so.

Neighbour indices and masks are piecewise-constant in the states, so - exactly as in the
reference's eager autograd (topk indices / comparisons carry no gradient) - they are treated
as constants of the backward pass.
"""
from typing import Optional, Sequence

import torch

from . import ops

def pack_params(params: Sequence[torch.Tensor]) -> torch.Tensor:
    """state_dict-ordered parameters -> one flat fp32 buffer (the C ABI's layout).  Repacked on every call:
    the copy is 51k floats, and no cache key sees in-place edits made through `p.data` (they do not bump
    `_version`), so a cache could silently serve stale weights."""
    return torch.cat([p.detach().reshape(-1) for p in params]).contiguous()


def _split_like(flat: torch.Tensor, params: Sequence[torch.Tensor]):
    out, off = [], 0
    for p in params:
        n = p.numel()
        out.append(flat[off:off + n].view_as(p))
        off += n
    return out


def _batched(t: torch.Tensor, last: int):
    """[N,last] -> ([1,N,last], True);  [B,N,last] -> (t, False)."""
    if t.dim() == 2:
        return t.unsqueeze(0), True
    assert t.dim() == 3 and t.shape[-1] == last
    return t, False


class CBFFunction(torch.autograd.Function):
    """h = CBF(states) on a fixed neighbour graph (cbf.py:21-45)."""

    # a differentiable REPAIRED forward writes the tile records of the backward (macbf_cbf_forward_record) when they
    # fit this budget - the backward then skips the recompute; larger problems recompute (macbf_cbf_backward)
    RECORD_BUDGET_BYTES = 2 << 30

    @staticmethod
    def apply_auto(states, obstacles, idx, semantics, *params):
        """apply(), with the recording decided by the caller's grad mode (inside forward() grad mode is always off)."""
        wants_grad = torch.is_grad_enabled() and (states.requires_grad or any(p.requires_grad for p in params))
        return CBFFunction.apply(states, obstacles, idx, semantics, wants_grad, *params)

    @staticmethod
    def forward(ctx, states, obstacles, idx, semantics, wants_grad, *params):
        flat = pack_params(params)
        B, N, K = idx.shape
        ctx.record = None
        if (semantics == ops.REPAIRED and wants_grad and B * N * K < 2 ** 31
                and ops.cbf_record_bytes(B, N, K) <= CBFFunction.RECORD_BUDGET_BYTES):
            h, ctx.record = ops.cbf_forward_record(states, idx, flat, obstacles)
        else:
            h = ops.cbf_forward(states, idx, flat, obstacles, semantics)
        ctx.save_for_backward(states, idx, flat, obstacles if obstacles is not None else states.new_empty(0))
        ctx.has_obs = obstacles is not None
        ctx.semantics = semantics
        ctx.nparams = len(params)
        ctx.param_shapes = [p.shape for p in params]
        return h

    @staticmethod
    def backward(ctx, grad_h):
        if ctx.semantics != ops.REPAIRED:
            raise RuntimeError("LITERAL semantics is forward-only (the reference's losses do not run; SURVEY App. A)")
        states, idx, flat, obs = ctx.saved_tensors
        obs = obs if ctx.has_obs else None
        if ctx.record is not None:
            gp, gs = ops.cbf_backward_record(idx, flat, grad_h.contiguous(), ctx.record, want_grad_states=ctx.needs_input_grad[0])
        else:
            gp, gs = ops.cbf_backward(states, idx, flat, grad_h.contiguous(), obs, want_grad_states=ctx.needs_input_grad[0])
        grads, off = [], 0
        for shp in ctx.param_shapes:
            n = int(torch.Size(shp).numel())
            grads.append(gp[off:off + n].view(shp))
            off += n
        return (gs, None, None, None, None, *grads)


class ControllerFunction(torch.autograd.Function):
    """u = Controller(states, goals) on a fixed neighbour graph (controller.py:31-63)."""

    @staticmethod
    def forward(ctx, states, goals, obstacles, idx, semantics, *params):
        flat = pack_params(params)
        u, _, _ = ops.controller_forward(states, goals, idx, flat, obstacles, semantics, want_next=False)
        ctx.save_for_backward(states, goals, idx, flat, obstacles if obstacles is not None else states.new_empty(0))
        ctx.has_obs = obstacles is not None
        ctx.semantics = semantics
        ctx.param_shapes = [p.shape for p in params]
        return u

    @staticmethod
    def backward(ctx, grad_u):
        if ctx.semantics != ops.REPAIRED:
            raise RuntimeError("LITERAL semantics is forward-only (the reference's losses do not run; SURVEY App. A)")
        states, goals, idx, flat, obs = ctx.saved_tensors
        obs = obs if ctx.has_obs else None
        gp, gs, gg = ops.controller_backward(states, goals, idx, flat, grad_u.contiguous(), None, obs,
                                             want_grad_states=ctx.needs_input_grad[0],
                                             want_grad_goals=ctx.needs_input_grad[1])
        grads, off = [], 0
        for shp in ctx.param_shapes:
            n = int(torch.Size(shp).numel())
            grads.append(gp[off:off + n].view(shp))
            off += n
        return (gs, gg, None, None, None, *grads)


class HingeLossFunction(torch.autograd.Function):
    """The 9 loss/accuracy scalars of core.py:89-184 for one batch of envs.
    Inputs h, h_next [B,N,K] (h_next optional), u [B,N,2] (optional); states/goals/idx are data.
    Output scalars [B,12] (layout: include/macbf_b200.h)."""

    @staticmethod
    def forward(ctx, h, h_next, u, states, goals, obstacles, idx):
        scalars, mask = ops.loss_forward(states, idx, h, h_next, u, goals, obstacles)
        ctx.save_for_backward(h, h_next if h_next is not None else h.new_empty(0),
                              u if u is not None else h.new_empty(0), states,
                              goals if goals is not None else h.new_empty(0), mask, scalars)
        ctx.flags = (h_next is not None, u is not None, goals is not None)
        ctx.mark_non_differentiable(mask)
        return scalars, mask

    @staticmethod
    def backward(ctx, grad_scalars, _grad_mask):
        h, h_next, u, states, goals, mask, scalars = ctx.saved_tensors
        has_next, has_u, has_goals = ctx.flags
        gh, ghn, gu, gst = ops.loss_backward(states, goals if has_goals else None, h, h_next if has_next else None,
                                             u if has_u else None, mask, scalars, grad_scalars.contiguous(),
                                             want_h=ctx.needs_input_grad[0],
                                             want_h_next=has_next and ctx.needs_input_grad[1],
                                             want_u=has_u and ctx.needs_input_grad[2],
                                             want_states=has_u and ctx.needs_input_grad[3])
        return gh, ghn, gu, gst, None, None, None


class RolloutFunction(torch.autograd.Function):
    """The closed loop of train.py:59-81 as ONE autograd node: forward = macbf_rollout_forward (stop test on the device, a
    single host read of the step counts at the end), backward = macbf_rollout_backward (BPTT through every controller
    call and Euler step of the saved trajectory; the reference does not detach it).
    -> (traj_states [T,B,N,4], traj_actions [T,B,N,2], steps_run [B]) with T = max over envs of steps_run; slots of an
    env beyond its own steps_run hold its continued simulation and receive no gradient."""

    @staticmethod
    def forward(ctx, states0, goals, obstacles, noise, max_steps, top_k, dist_stop, *params):
        flat = pack_params(params)
        ts, ta, steps = ops.rollout_forward(states0, goals, flat, max_steps, obstacles, noise, top_k=top_k,
                                            semantics=ops.REPAIRED, dist_stop=dist_stop)
        T = int(steps.max().item())          # the one host synchronisation of the rollout
        ctx.traj, ctx.steps, ctx.T = ts, steps, T
        ctx.save_for_backward(goals, flat, obstacles if obstacles is not None else states0.new_empty(0))
        ctx.has_obs = obstacles is not None
        ctx.top_k = top_k
        ctx.param_shapes = [p.shape for p in params]
        ctx.mark_non_differentiable(steps)
        return ts[:T], ta[:T], steps

    @staticmethod
    def backward(ctx, g_ts, g_ta, _g_steps):
        goals, flat, obs = ctx.saved_tensors
        obs = obs if ctx.has_obs else None
        T, steps = ctx.T, ctx.steps
        live = (torch.arange(T, device=steps.device)[:, None] < steps[None, :]).to(torch.float32)[:, :, None, None]
        g_ta = (g_ta * live).contiguous() if g_ta is not None else torch.zeros_like(ctx.traj[:T, ..., :2])
        g_ts = (g_ts * live).contiguous() if g_ts is not None else None
        gp, gs0, gg = ops.rollout_backward(ctx.traj, goals, flat, g_ta, g_ts, obs, top_k=ctx.top_k,
                                           want_grad_states0=ctx.needs_input_grad[0],
                                           want_grad_goals=ctx.needs_input_grad[1])
        ctx.traj = None
        grads, off = [], 0
        for shp in ctx.param_shapes:
            n = int(torch.Size(shp).numel())
            grads.append(gp[off:off + n].view(shp))
            off += n
        return (gs0, gg, None, None, None, None, None, *grads)
