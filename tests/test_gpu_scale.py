"""GPU: BASELINE.json's larger configurations through size-independent properties and sampled oracle checks.

C3  4096 agents + 384 obstacle points (16 circles + 16 rectangles), full training step: one env is
    small enough for the dense CPU oracle (a few seconds), so it is compared directly.
C4  16384-agent envs: neighbour sets of sampled rows against the row-chunked oracle, structure checks.
C5  65536 agents, K in {8, 32, 64, 128} with the sensing radius scaled to the neighbour count.
"""
import math

import numpy as np
import pytest
import torch

from helpers import RTOL, assert_close, assert_scalars_close, dev, oracle_step_on_device_graph, rel_err

GTOL = 5e-5   # gradients: sums over ~10^4..10^6 edges of fp32 products (tests/test_gpu_backward.py)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from macbf_b200 import ops as _ops
    return _ops


def _params():
    from oracle import macbf_oracle as O
    p_cbf, p_ctrl = O.init_params(21)
    return p_cbf, p_ctrl, dev(O.flatten_params(p_cbf, O.CBF_KEYS).numpy()), dev(O.flatten_params(p_ctrl, O.CTRL_KEYS).numpy())


def test_c3_training_step_with_obstacles(ops):
    from macbf_b200 import step
    from oracle import macbf_oracle as O
    N = 4096
    s, g = O.synthetic_swarm(1, N, seed=1234 + 3000)
    obs = O.synthetic_obstacles(1, N, 16, 16, seed=1234 + 3000)
    assert obs.shape == (1, 384, 2)
    p_cbf, p_ctrl, cf, tf = _params()
    ts = step.TrainStep(1, N, 384, 12, "cuda")
    scal, grad = ts.forward_backward(s.cuda(), g.cuda(), cf, tf, obs.cuda())
    torch.cuda.synchronize()
    # losses and gradients are compared on the device's S' graph (rows that differ from the oracle's own top-k are first
    # proven to be near-ties), so nothing below is conditional
    ref, frac_same = oracle_step_on_device_graph(p_cbf, p_ctrl, s[0], g[0], obs[0], 12,
                                                 ts.buf.idx_next[0].cpu().numpy(), min_same=0.995)
    idx_ref = O.neighbor_indices(s[0], 12, obs[0], row_chunk=512).numpy()
    dec = O.decidable_rows(s[0], 12, obs[0]).numpy()
    assert np.array_equal(ts.buf.idx[0].cpu().numpy()[dec], idx_ref[dec]) and dec.mean() > 0.999
    assert (ts.buf.idx[0].cpu().numpy() >= N).any()                 # some neighbours are obstacle points
    assert_close(ts.buf.u[0].cpu().numpy(), ref["u"].numpy(), RTOL, "u")
    assert_close(ts.buf.h[0].cpu().numpy(), ref["h"][:, 0, :].numpy(), RTOL, "h")
    assert_close(ts.buf.h_next[0].cpu().numpy(), ref["h_next"][:, 0, :].numpy(), 2 * RTOL, "h_next")
    assert np.array_equal(ts.buf.ttc_mask[0].cpu().numpy().astype(bool), ref["ttc_mask"][:, :, 0].numpy())
    sc = scal[0].cpu().numpy()
    assert_scalars_close(sc, ref["scalars"].numpy(), float(ref["h"].abs().max()), sc[10], sc[11], "C3")
    assert rel_err(grad[:ops.CBF_PARAMS].cpu().numpy(), ref["grad_cbf"].numpy()) <= GTOL
    assert rel_err(grad[ops.CBF_PARAMS:].cpu().numpy(), ref["grad_ctrl"].numpy()) <= GTOL
    print(f"C3: {frac_same:.5f} of the S' rows identical to the oracle's own top-k")


def test_c2_training_step_gradient_all_64_envs(ops):
    """BASELINE configs[1] shape in TRAINING mode (the path bench.py's `train` block times): 64 envs x 1024 agents,
    gradient of the env-mean objective against the oracle's autograd summed over all 64 envs, plus the 9 scalars of
    every env (the CPU side costs ~0.3 s per env)."""
    from macbf_b200 import step
    from oracle import macbf_oracle as O
    B, N = 64, 1024
    s, g = O.synthetic_swarm(B, N, seed=1234 + 2000)
    p_cbf, p_ctrl, cf, tf = _params()
    ts = step.TrainStep(B, N, 0, 12, "cuda")
    scal, grad = ts.forward_backward(s.cuda(), g.cuda(), cf, tf)
    torch.cuda.synchronize()
    idx_next = ts.buf.idx_next.cpu().numpy()
    sc = scal.cpu().numpy()
    want_c = torch.zeros(ops.CBF_PARAMS, dtype=torch.float64)
    want_t = torch.zeros(ops.CTRL_PARAMS, dtype=torch.float64)
    worst_same = 1.0
    for b in range(B):
        ref, same = oracle_step_on_device_graph(p_cbf, p_ctrl, s[b], g[b], None, 12, idx_next[b])
        worst_same = min(worst_same, same)
        want_c += ref["grad_cbf"].double() / B
        want_t += ref["grad_ctrl"].double() / B
        assert_scalars_close(sc[b], ref["scalars"].numpy(), float(ref["h"].abs().max()), sc[b, 10], sc[b, 11], f"env {b}")
    assert rel_err(grad[:ops.CBF_PARAMS].cpu().numpy(), want_c.numpy()) <= GTOL
    assert rel_err(grad[ops.CBF_PARAMS:].cpu().numpy(), want_t.numpy()) <= GTOL
    print(f"C2 train: worst env has {worst_same:.5f} of its S' rows identical to the oracle's own top-k")


def _sampled_rows_values(ops, s, g, b, rows, k, cf, tf, p_cbf, p_ctrl, h_dev, u_dev, idx_dev, R=1.0):
    """h and u of sampled query rows against the row-chunked oracle (dense N x N never has to fit): the oracle forms
    the sampled rows' full candidate lists, selects their top-k itself and runs the two nets on them."""
    from oracle import macbf_oracle as O
    r = torch.from_numpy(rows)
    d = s[b][r][:, None, :2] - s[b][None, :, :2]
    keys = O.knn_keys(d)
    v, ref_idx = torch.topk(-keys, k + 1)
    v = -v
    dec = ((v[:, 1:].view(torch.int32) - v[:, :-1].view(torch.int32)) >= 2).all(dim=1).numpy()
    assert dec.mean() > 0.95
    assert np.array_equal(idx_dev[rows][dec], ref_idx[:, :k].numpy()[dec])
    h_ref = O.cbf_forward(p_cbf, s[b], O.REPAIRED, None, k, R, rows=r)[:, 0, :].numpy()
    u_ref = O.controller_forward(p_ctrl, s[b], g[b], O.REPAIRED, None, k, R, rows=r).numpy()
    assert_close(h_dev[rows][dec], h_ref[dec], RTOL, "h (sampled rows)")
    assert_close(u_dev[rows][dec], u_ref[dec], RTOL, "u (sampled rows)")


def _structure_checks(idx, N, C, K):
    assert idx.min() >= 0 and idx.max() < C
    assert (idx[:, 0] == np.arange(N)).all()                         # slot 0 is the agent itself
    srt = np.sort(idx, axis=1)
    assert (srt[:, 1:] != srt[:, :-1]).all()                          # no duplicates


def test_c4_large_envs_sampled_rows(ops):
    from oracle import macbf_oracle as O
    B, N = 3, 16384
    s, g = O.synthetic_swarm(B, N, seed=1234 + 4000)
    idx = ops.neighbor_topk(s.cuda(), 12).cpu().numpy()
    rows = np.random.RandomState(0).choice(N, 512, replace=False)
    for b in range(B):
        _structure_checks(idx[b], N, N, 12)
        d = s[b, rows, None, :2] - s[b, None, :, :2]
        ref = torch.topk(-O.knn_keys(d), 12)[1].numpy()
        assert np.array_equal(idx[b][rows], ref)
    _, _, cf, tf = _params()
    buf = ops.StepBuffers(B, N, 0, 12, "cuda")
    ops.step_forward(s.cuda(), g.cuda(), cf, tf, buf)
    torch.cuda.synchronize()
    sc = buf.scalars.cpu().numpy()
    assert np.isfinite(sc).all() and (sc[:, 10] + sc[:, 11] == N * 12).all()
    # h is exactly zero where (and only where) the sensing-radius mask is off: check against geometry
    sd = s[0]
    j = torch.from_numpy(idx[0].astype(np.int64))
    dxy = sd[:, None, :2] - sd[j][:, :, :2]
    dist = torch.sqrt((dxy[..., 0] ** 2 + 1e-4) + (dxy[..., 1] ** 2 + 1e-4))
    masked = (dist > 1.0).numpy()
    h0 = buf.h[0].cpu().numpy()
    assert (h0[masked] == 0).all()
    p_cbf, p_ctrl, _, _ = _params()
    for b in range(B):
        _sampled_rows_values(ops, s, g, b, rows, 12, cf, tf, p_cbf, p_ctrl, buf.h[b].cpu().numpy(), buf.u[b].cpu().numpy(),
                             idx[b])


@pytest.mark.parametrize("k", [8, 32, 64, 128])
def test_c5_density_sweep_65536(ops, k):
    from oracle import macbf_oracle as O
    N = 65536
    s, g = O.synthetic_swarm(1, N, seed=1234 + 5000 + k)
    idx = ops.neighbor_topk(s.cuda(), k)[0].cpu().numpy()
    _structure_checks(idx, N, N, k)
    rows = np.random.RandomState(k).choice(N, 256, replace=False)
    d = s[0, rows, None, :2] - s[0, None, :, :2]
    keys = O.knn_keys(d)
    vals, ref = torch.topk(-keys, k + 1)
    vals = -vals
    gap_ok = ((vals[:, 1:].view(torch.int32) - vals[:, :-1].view(torch.int32)) >= 2).all(dim=1).numpy()
    assert np.array_equal(idx[rows][gap_ok], ref[:, :k].numpy()[gap_ok]) and gap_ok.mean() > 0.95
    # keys along every row are non-decreasing (sortedness)
    jj = torch.from_numpy(idx[rows].astype(np.int64))
    kk = torch.gather(keys, 1, jj)
    assert (kk[:, 1:] >= kk[:, :-1]).all()
    # radius scaled to the neighbour count: R = sqrt(K / (8 pi)); CBF + controller run at that K
    R = math.sqrt(k / (8 * math.pi))
    _, _, cf, tf = _params()
    idx_d = torch.from_numpy(idx).cuda().unsqueeze(0)
    h = ops.cbf_forward(s.cuda(), idx_d, cf, obs_radius=R)
    u, sn, pool = ops.controller_forward(s.cuda(), g.cuda(), idx_d, tf, obs_radius=R)
    assert torch.isfinite(h).all() and torch.isfinite(u).all() and torch.isfinite(sn).all()
    p_cbf, p_ctrl, _, _ = _params()
    _sampled_rows_values(ops, s, g, 0, rows[:128], k, cf, tf, p_cbf, p_ctrl, h[0].cpu().numpy(), u[0].cpu().numpy(), idx, R)
    frac_in_range = float((h != 0).float().mean())
    assert 0.2 < frac_in_range < 0.995                                # R holds K agents on average: most, not all, of the K nearest are inside


@pytest.mark.parametrize("N,k,n_obs,nparts", [(65536, 12, 0, 8), (5000, 12, 2, 3), (700, 64, 0, 4), (200, 12, 0, 2), (9, 12, 0, 2), (11, 12, 0, 4)])
def test_row_sharded_graph_parts_assemble_the_full_graph(ops, N, k, n_obs, nparts):
    """one large env row-sharded over ranks - every part fills a disjoint set of rows and the sum of the
    zero-initialised partial index arrays (what the all-reduce computes) is the full graph, bit for bit."""
    import oracle.macbf_oracle as O
    s, _ = O.synthetic_swarm(1, N, 5)
    obs = O.synthetic_obstacles(1, N, n_obs, 0, 6).cuda() if n_obs else None
    s = s.cuda()
    full = ops.neighbor_topk(s, k, obs)
    total = torch.zeros_like(full)
    owners = torch.zeros(full.shape[:2], dtype=torch.int32, device="cuda")
    for p in range(nparts):
        marked = torch.full_like(full, -1)
        ops.neighbor_topk_part(s, k, p, nparts, obs, out=marked)
        rows = (marked >= 0).all(dim=2)
        assert torch.equal((marked >= 0).any(dim=2), rows)            # a row is written completely or not at all
        owners += rows.int()
        total += ops.neighbor_topk_part(s, k, p, nparts, obs)
    assert int(owners.min()) == 1 and int(owners.max()) == 1          # disjoint cover (also when N + M <= k: no top-k)
    assert torch.equal(total, full)
