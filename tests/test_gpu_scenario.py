"""GPU: macbf_generate_scenario (core.py:45-71 on the device) against the oracle's SEQUENTIAL restatement of the same
dart-throwing process on the same counter-based stream (identical bits), plus the properties the reference's
generator guarantees."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
THRES = 0.07


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from macbf_b200 import ops as _ops
    return _ops


def _min_pair_dist(p):
    d = np.linalg.norm(p[:, None, :].astype(np.float64) - p[None, :, :].astype(np.float64), axis=2)
    np.fill_diagonal(d, np.inf)
    return d.min()


@pytest.mark.parametrize("N,B,seed", [(64, 3, 7), (300, 2, 123456789012345), (5, 2, 1)])
def test_device_scenario_equals_sequential_process(ops, N, B, seed):
    import oracle.macbf_oracle as O
    s, g = ops.generate_scenario(seed, B, N, THRES)
    s, g = s.cpu().numpy(), g.cpu().numpy()
    for b in range(B):
        want = O.generate_data_stream(seed, b, N, THRES)
        assert want is not None
        assert np.array_equal(s[b], want[0]), f"env {b}: start positions differ"
        assert np.array_equal(g[b], want[1]), f"env {b}: goals differ"


def test_device_scenario_properties_large(ops):
    N, B = 4096, 4                                   # the reference's numpy loop needs 1.6 s for ONE such env
    s, g = ops.generate_scenario(99, B, N, THRES)
    s2, g2 = ops.generate_scenario(99, B, N, THRES)
    assert torch.equal(s, s2) and torch.equal(g, g2)                         # deterministic in (seed, env)
    s3, _ = ops.generate_scenario(100, B, N, THRES)
    assert not torch.equal(s, s3)
    s, g = s.cpu().numpy(), g.cpu().numpy()
    side = np.sqrt(N / 8.0)
    assert np.all(s[:, :, 2:] == 0.0)                                        # core.py:68-69: zero initial velocity
    assert s[:, :, :2].min() >= 0.0 and s[:, :, :2].max() < side
    assert np.abs(g - s[:, :, :2]).max() <= 0.5 + 1e-6                       # core.py:59
    for b in range(B):
        assert _min_pair_dist(s[b, :, :2]) > THRES
        assert _min_pair_dist(g[b]) > THRES
        assert np.linalg.norm(s[b, :, :2], axis=1).min() > THRES             # the origin quirk of core.py:47-54
    assert not np.array_equal(s[0], s[1])                                    # envs are independent
    # uniform coverage: every quadrant holds about a quarter of the agents
    q = (s[0, :, 0] > side / 2).astype(int) * 2 + (s[0, :, 1] > side / 2).astype(int)
    assert np.all(np.abs(np.bincount(q, minlength=4) / N - 0.25) < 0.04)


def test_device_scenario_feeds_the_step(ops):
    import oracle.macbf_oracle as O
    s, g = ops.generate_scenario(3, 2, 256, THRES)
    p_cbf, p_ctrl = O.init_params(1)
    cf, tf = O.flatten_params(p_cbf, O.CBF_KEYS).cuda(), O.flatten_params(p_ctrl, O.CTRL_KEYS).cuda()
    buf = ops.StepBuffers(2, 256, 0, 12, "cuda")
    ops.step_forward(s, g, cf, tf, buf)
    want = O.full_step(p_cbf, p_ctrl, s[1].cpu(), g[1].cpu())
    assert np.array_equal(buf.idx[1].cpu().numpy(), O.neighbor_indices(s[1].cpu(), 12).numpy())
    np.testing.assert_allclose(buf.scalars[1, :9].cpu().numpy(), want["scalars"].numpy(), rtol=2e-5, atol=1e-6)


def test_device_obstacle_point_sets_equal_reference(ops):
    """macbf_generate_obstacles (core.py:7-42 on the device) against the point sets the unmodified reference produced
    (tests/golden/ref_obstacles.npz): straight segments bit-identical in float64; circle points within 2 ulp in float64
    (device cos/sin vs libm) and identical once cast to the float32 the kernels consume."""
    import os
    from helpers import OBSTACLE_SPECS
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    z = np.load(os.path.join(root, "tests", "golden", "ref_obstacles.npz"))
    got = ops.generate_obstacles(OBSTACLE_SPECS, 12).cpu().numpy().reshape(len(OBSTACLE_SPECS), 12, 2)
    for i, (kind, _, _) in enumerate(OBSTACLE_SPECS):
        want = z[f"p{i}"]
        if kind == "rectangle":
            assert np.array_equal(got[i], want), f"rectangle {i}"
        else:
            assert np.abs(got[i] - want).max() <= 2 * np.spacing(np.abs(want).max()), f"circle {i}"
        assert np.array_equal(got[i].astype(np.float32), want.astype(np.float32)), f"obstacle {i} as float32"
    g7 = ops.generate_obstacles([OBSTACLE_SPECS[0]], 7).cpu().numpy()
    assert np.array_equal(g7.astype(np.float32), z["circle_num7"].astype(np.float32))
    g30 = ops.generate_obstacles([OBSTACLE_SPECS[1]], 30).cpu().numpy()
    assert np.array_equal(g30, z["rect_num30"])
    # and as candidate columns of the neighbour search: same graph as with the host-made point set
    import oracle.macbf_oracle as O
    s, _ = O.synthetic_swarm(1, 64, 3)
    pts = ops.generate_obstacles(OBSTACLE_SPECS[:2], 12).to(torch.float32).unsqueeze(0)
    host = torch.from_numpy(np.concatenate([z["p0"], z["p1"]]).astype(np.float32)).unsqueeze(0).cuda()
    assert torch.equal(ops.neighbor_topk(s.cuda(), 12, pts), ops.neighbor_topk(s.cuda(), 12, host))
