"""GPU: the tensor-core backward kernels (cbf_bwd_tc.cu, ctrl_edge_bwd_tc.cu, ctrl_node_bwd_tc.cu) against torch autograd
of the oracle over tile geometries the golden fixtures do not reach: K from 1 to 128, ragged last tiles, obstacles,
multi-chunk records; run-to-run determinism; saved winners vs recompute."""
import os

import pytest
import torch

from helpers import dev

pytestmark = pytest.mark.gpu
GTOL = 5e-5   # scale-relative, as in test_gpu_backward.py


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from macbf_b200 import ops as _ops
    return _ops


def _rel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _run(ops, s, g, obs, idx, cf, tf, gh, gu, gs):
    gp, gst = ops.cbf_backward(s, idx, cf, gh, obstacles=obs, want_grad_states=True)
    cp, cst, cg = ops.controller_backward(s, g, idx, tf, gu, gs, obstacles=obs, want_grad_states=True, want_grad_goals=True)
    torch.cuda.synchronize()
    return [x.clone() for x in (gp, gst, cp, cst, cg)]


def _case(ops, B, N, K, n_obs, seed):
    import oracle.macbf_oracle as O
    s, g = O.synthetic_swarm(B, N, seed)
    obs = dev(O.synthetic_obstacles(B, N, n_obs, 0, seed + 1)) if n_obs else None
    p_cbf, p_ctrl = O.init_params(seed)
    cf, tf = dev(O.flatten_params(p_cbf, O.CBF_KEYS)), dev(O.flatten_params(p_ctrl, O.CTRL_KEYS))
    s, g = dev(s), dev(g)
    idx = ops.neighbor_topk(s, K, obs)
    Ke = idx.shape[2]
    gen = torch.Generator(device="cuda").manual_seed(seed)
    gh = torch.randn(B, N, Ke, device="cuda", generator=gen) * (torch.rand(B, N, Ke, device="cuda", generator=gen) < 0.4)
    gu = torch.randn(B, N, 2, device="cuda", generator=gen)
    gs = torch.randn(B, N, 4, device="cuda", generator=gen)
    return s, g, obs, idx, cf, tf, gh, gu, gs


def _with_env(env, fn):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        return fn()
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


NAMES = ["cbf params", "cbf grad_states", "ctrl params", "ctrl grad_states", "ctrl grad_goals"]


def test_multi_chunk_records_match_single_chunk(ops):
    args = _case(ops, 4, 257, 12, 0, 5)          # 1028 agents / 10 per tile = 103 tiles
    one = _run(ops, *args)
    many = _with_env({"MACBF_BWD_CHUNK_TILES": "7"}, lambda: _run(ops, *args))
    for name, a, b in zip(NAMES, many, one):
        # same kernels, different partition of the tiles over CTAs: only the fp32 summation order changes
        assert _rel(a, b) < 2e-6, f"{name} differs by {_rel(a, b):.2e}"


def test_backward_is_run_to_run_deterministic_for_parameters(ops):
    args = _case(ops, 2, 300, 12, 0, 9)
    a = _run(ops, *args)
    b = _run(ops, *args)
    assert torch.equal(a[0], b[0]) and torch.equal(a[2], b[2])   # parameter gradients: fixed reduction order


def test_saved_pool_and_winners_give_the_same_gradients_as_the_recompute(ops):
    """macbf_step_forward_ex(MACBF_STEP_SAVE_WINNERS) + macbf_controller_backward_saved skip the edge-net recompute:
    the gradients must be bit-identical to macbf_controller_backward (same kernels, same inputs)."""
    from macbf_b200 import lib
    import oracle.macbf_oracle as O
    B, N, K = 3, 333, 12
    s, g, obs, idx, cf, tf, gh, gu, gs = _case(ops, B, N, K, 2, 4)
    M = obs.shape[1]
    buf = ops.StepBuffers(B, N, M, K, "cuda")
    ops.step_forward(s, g, cf, tf, buf, obs, save_winners=True)
    assert torch.equal(buf.idx, idx)
    want, want_s, want_g = ops.controller_backward(s, g, idx, tf, gu, gs, obstacles=obs, want_grad_states=True, want_grad_goals=True)
    L = lib.load()
    grad = torch.zeros_like(want); gst = torch.zeros_like(want_s); gg = torch.zeros_like(want_g)
    nb = L.macbf_controller_backward_scratch_bytes(B, N, K)
    scratch = torch.empty((nb,), dtype=torch.uint8, device="cuda")
    ws = buf.workspace.data_ptr()
    lib.check(L.macbf_controller_backward_saved(s.data_ptr(), g.data_ptr(), obs.data_ptr(), idx.data_ptr(), tf.data_ptr(),
                                                gu.data_ptr(), gs.data_ptr(), B, N, M, K, ops.OBS_RADIUS, ops.TIME_STEP, ws,
                                                ws + buf.winners_offset, grad.data_ptr(), gst.data_ptr(), gg.data_ptr(),
                                                scratch.data_ptr(), nb, None), "macbf_controller_backward_saved")
    torch.cuda.synchronize()
    assert torch.equal(grad, want)
    assert torch.equal(gg, want_g)
    assert _rel(gst, want_s) < 1e-6          # d/dstates goes through fp32 atomics (order varies)
    assert float(grad.abs().max()) > 0


@pytest.mark.parametrize("K,N,n_obs", [(1, 150, 0), (3, 150, 0), (5, 333, 2), (12, 150, 0), (20, 150, 3), (64, 150, 0), (128, 200, 0)])
def test_tensor_core_backward_matches_oracle_autograd(ops, K, N, n_obs):
    """The same tile geometries against torch autograd of the ORACLE (not only against the FFMA twin): for random
    cotangents, d/d(params, states, goals) of sum(h * gh) and of sum(u * gu) + sum(s' * gs)
    (autograd of cbf.py:21-45 and controller.py:31-63 + train.py:70 on a fixed neighbour graph)."""
    import oracle.macbf_oracle as O
    B = 2
    s, g, obs, idx, cf, tf, gh, gu, gs = _case(ops, B, N, K, n_obs, 40 + K)
    tc = _run(ops, s, g, obs, idx, cf, tf, gh, gu, gs)
    p_cbf, p_ctrl = O.init_params(40 + K)
    pc = {n: t.clone().requires_grad_(True) for n, t in p_cbf.items()}
    pt = {n: t.clone().requires_grad_(True) for n, t in p_ctrl.items()}
    want = [None, torch.zeros(B, N, 4), None, torch.zeros(B, N, 4), torch.zeros(B, N, 2)]
    for b in range(B):
        sb = s[b].cpu().clone().requires_grad_(True)
        gb = g[b].cpu().clone().requires_grad_(True)
        ob = obs[b].cpu() if obs is not None else None
        ind = idx[b].cpu().long()
        assert torch.equal(ind, O.neighbor_indices(sb.detach(), K, ob))
        h = O.cbf_forward(pc, sb, O.REPAIRED, ob, K)
        (h[:, 0, :] * gh[b].cpu()).sum().backward()
        want[1][b] = sb.grad.clone()
        sb.grad = None
        u = O.controller_forward(pt, sb, gb, O.REPAIRED, ob, K)
        ((u * gu[b].cpu()).sum() + (O.euler_step(sb, u) * gs[b].cpu()).sum()).backward()
        want[3][b], want[4][b] = sb.grad.clone(), gb.grad.clone()
    want[0] = torch.cat([pc[n].grad.reshape(-1) for n in O.CBF_KEYS])
    want[2] = torch.cat([pt[n].grad.reshape(-1) for n in O.CTRL_KEYS])
    for name, a, b in zip(NAMES, tc, want):
        assert _rel(a.cpu(), b) < GTOL, f"K={K}: {name} differs from oracle autograd by {_rel(a.cpu(), b):.2e}"


@pytest.mark.parametrize("B,N,K,n_obs", [(2, 200, 12, 0), (1, 96, 5, 2), (2, 150, 64, 0), (1, 200, 128, 0), (3, 7, 12, 0), (1, 1000, 1, 0)])
def test_recording_forward_and_its_backward_equal_the_recompute_path(ops, B, N, K, n_obs):
    """macbf_cbf_forward_record + macbf_cbf_backward_record (the training step's pair: the forward writes the tile records,
    the backward starts at the delta chain) against macbf_cbf_forward + macbf_cbf_backward (forward, then recompute):
    same h bit for bit (a row's value does not depend on the tile it sits in), same gradients up to the summation
    order of dW4 / db4 (which move from the dump kernel to the delta-chain kernel)."""
    s, g, obs, idx, cf, tf, gh, gu, gs = _case(ops, B, N, K, n_obs, 21)
    h_plain = ops.cbf_forward(s, idx, cf, obs)
    h_rec, rec = ops.cbf_forward_record(s, idx, cf, obs)
    assert rec.numel() == ops.cbf_record_bytes(B, N, idx.shape[2])
    assert torch.equal(h_plain, h_rec)
    gp1, gs1 = ops.cbf_backward(s, idx, cf, gh, obstacles=obs, want_grad_states=True)
    gp2, gs2 = ops.cbf_backward_record(idx, cf, gh, rec, want_grad_states=True)
    assert _rel(gp2, gp1) <= 2e-6, f"grad_params record vs recompute: {_rel(gp2, gp1):.2e}"
    assert _rel(gs2, gs1) <= 2e-6, f"grad_states record vs recompute: {_rel(gs2, gs1):.2e}"
    # W4 / b4 (the part computed by different code): tight on their own scale
    from macbf_b200.ops import CBF_PARAMS
    w4 = slice(CBF_PARAMS - 65, CBF_PARAMS)
    assert _rel(gp2[w4], gp1[w4]) <= 2e-6, f"dW4/db4 record vs recompute: {_rel(gp2[w4], gp1[w4]):.2e}"


def test_train_step_record_and_recompute_modes_agree(ops):
    """TrainStep(record_forward=True) (default when the records fit) against record_forward=False on the same inputs:
    same scalars bit for bit, same flat gradient to summation order."""
    import oracle.macbf_oracle as O
    from macbf_b200.step import TrainStep
    B, N = 3, 300
    s, g = O.synthetic_swarm(B, N, 4)
    obs = dev(O.synthetic_obstacles(B, N, 2, 0, 5))
    p_cbf, p_ctrl = O.init_params(4)
    cf, tf = dev(O.flatten_params(p_cbf, O.CBF_KEYS)), dev(O.flatten_params(p_ctrl, O.CTRL_KEYS))
    out = []
    for rec in (True, False):
        ts = TrainStep(B, N, obs.shape[1], 12, "cuda", record_forward=rec)
        assert ts.record_forward is rec
        sc, gr = ts.forward_backward(dev(s), dev(g), cf, tf, obs)
        out.append((sc.clone(), gr.clone()))
    assert TrainStep(B, N, obs.shape[1], 12, "cuda").record_forward is True
    assert torch.equal(out[0][0], out[1][0])
    assert _rel(out[0][1], out[1][1]) <= 2e-6, f"{_rel(out[0][1], out[1][1]):.2e}"


def test_record_entry_points_reject_small_buffers(ops):
    from macbf_b200 import lib
    L = lib.load()
    s, g, obs, idx, cf, tf, gh, gu, gs = _case(ops, 1, 64, 12, 0, 3)
    h = torch.empty(1, 64, 12, device="cuda")
    small = torch.empty(4096, dtype=torch.uint8, device="cuda")
    rc = L.macbf_cbf_forward_record(s.data_ptr(), None, idx.data_ptr(), cf.data_ptr(), 1, 64, 0, 12, 0.07, 1.0, h.data_ptr(),
                                    small.data_ptr(), small.numel(), None)
    assert rc == 1 and b"record buffer too small" in L.macbf_last_error()
    gp = torch.zeros(17089, device="cuda")
    rc = L.macbf_cbf_backward_record(idx.data_ptr(), cf.data_ptr(), gh.data_ptr(), 1, 64, 12, small.data_ptr(), small.numel(),
                                     gp.data_ptr(), None, small.data_ptr(), small.numel(), None)
    assert rc == 1 and b"too small" in L.macbf_last_error()
