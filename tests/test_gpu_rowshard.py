"""GPU: one large env row-sharded over G logical ranks (macbf_step_forward_rows_a/_b/_c) against the
unsharded fused step: every per-row output bit for bit, the scalars to 1e-6 (the action term's fp32 additions are
grouped differently).  The G ranks run one after the other on one GPU; the multi-GPU run of the same schedule over
NCCL is scripts/r2_rowshard.sh."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from macbf_b200 import ops as _ops
    return _ops


@pytest.mark.parametrize("N,k,n_obs,nparts", [(8192, 12, 0, 4), (5000, 12, 2, 3), (3072, 40, 0, 2), (4096, 64, 0, 8), (2048, 12, 0, 1)])
def test_row_sharded_step_equals_unsharded(ops, N, k, n_obs, nparts):
    from macbf_b200 import rowshard
    from oracle import macbf_oracle as O
    s, g = O.synthetic_swarm(1, N, 77 + N)
    obs = O.synthetic_obstacles(1, N, n_obs, 0, 5).cuda() if n_obs else None
    M = 0 if obs is None else obs.shape[1]
    p_cbf, p_ctrl = O.init_params(3)
    cf, tf = O.flatten_params(p_cbf, O.CBF_KEYS).cuda(), O.flatten_params(p_ctrl, O.CTRL_KEYS).cuda()
    s, g = s.cuda(), g.cuda()
    ref = ops.StepBuffers(1, N, M, k, "cuda")
    ops.step_forward(s, g, cf, tf, ref, obs)
    step = rowshard.RowShardedStep(N, M, k, "cuda")
    for t in (step.buf.idx, step.buf.idx_next):
        t.fill_(-1)
    for t in (step.buf.u, step.buf.s_next, step.buf.h, step.buf.h_next):
        t.fill_(float("nan"))
    got = rowshard.forward_parts_on_one_gpu(step, nparts, s, g, cf, tf, obs)
    torch.cuda.synchronize()
    for name in ("idx", "idx_next", "u", "s_next", "h", "h_next", "ttc_mask"):
        assert torch.equal(getattr(got, name), getattr(ref, name)), name
    a, b = got.scalars.cpu().numpy(), ref.scalars.cpu().numpy()
    assert np.array_equal(a[:, [10, 11]], b[:, [10, 11]])
    assert np.abs(a - b).max() <= 1e-6 * max(1.0, np.abs(b).max())
    assert np.array_equal(a[:, :8], b[:, :8])          # the hinge sums keep their chunks: bit-identical


def test_row_range_validation(ops):
    from macbf_b200 import lib, rowshard
    step = rowshard.RowShardedStep(4096, 0, 12, "cuda")
    s = torch.zeros(1, 4096, 4, device="cuda")
    g = torch.zeros(1, 4096, 2, device="cuda")
    tf = torch.zeros(ops.CTRL_PARAMS, device="cuda")
    with pytest.raises(lib.MacbfError, match="multiples of 1024"):
        step.phase_a(s, g, tf, None, 100, 2048)
    with pytest.raises(lib.MacbfError, match="outside"):
        step.phase_a(s, g, tf, None, 0, 8192)
