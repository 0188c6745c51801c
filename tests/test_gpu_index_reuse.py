"""GPU parity of the opt-in index-reuse variant of the derivative loss (h' = CBF(S') on the neighbour
sets of S, MACBF_STEP_REUSE_INDICES) against the oracle's same extension, forward and gradients."""
import numpy as np
import pytest
import torch

from helpers import RTOL, assert_close, dev

pytestmark = pytest.mark.gpu
GTOL = 5e-5


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from macbf_b200 import ops as _ops
    return _ops


@pytest.mark.parametrize("N,n_obs", [(96, 2), (400, 0)])
def test_index_reuse_step_matches_oracle(ops, N, n_obs):
    import oracle.macbf_oracle as O
    from macbf_b200.step import TrainStep
    s, g = O.synthetic_swarm(1, N, 31, vel_std=0.5)
    obs = O.synthetic_obstacles(1, N, n_obs, 0, 8) if n_obs else None
    p_cbf, p_ctrl = O.init_params(3)
    want = O.full_step(p_cbf, p_ctrl, s[0], g[0], None if obs is None else obs[0], with_grad=True, reuse_indices=True)
    plain = O.full_step(p_cbf, p_ctrl, s[0], g[0], None if obs is None else obs[0], reuse_indices=False)
    cf, tf = dev(O.flatten_params(p_cbf, O.CBF_KEYS)), dev(O.flatten_params(p_ctrl, O.CTRL_KEYS))
    M = 0 if obs is None else obs.shape[1]
    ts = TrainStep(1, N, M, 12, "cuda")
    scal, grad = ts.forward_backward(dev(s), dev(g), cf, tf, None if obs is None else dev(obs), reuse_indices=True)
    b = ts.buf
    assert torch.equal(b.idx_next, b.idx)
    assert_close(b.h_next[0].cpu().numpy(), want["h_next"][:, 0, :].numpy(), RTOL, "h_next (reused indices)")
    assert_close(scal[0, :9].cpu().numpy(), want["scalars"].numpy(), RTOL, "scalars")
    gc = grad[:cf.numel()].cpu().numpy()
    gt = grad[cf.numel():].cpu().numpy()
    assert_close(gc, want["grad_cbf"].numpy(), GTOL, "grad_cbf")
    assert_close(gt, want["grad_ctrl"].numpy(), GTOL, "grad_ctrl")
    # the variant really is a different computation on these inputs (fast agents change their neighbour sets)
    assert not np.allclose(want["h_next"].numpy(), plain["h_next"].numpy())
    # and the default path still recomputes the graph on S'
    ops.step_forward(dev(s), dev(g), cf, tf, b, None if obs is None else dev(obs))
    assert not torch.equal(b.idx_next, b.idx)
    assert_close(b.h_next[0].cpu().numpy(), plain["h_next"][:, 0, :].numpy(), RTOL, "h_next (fresh top-k)")
