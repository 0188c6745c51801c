"""CPU: the C-ABI library loads and exports exactly what include/macbf_b200.h declares; argument
validation works without touching a GPU; the drop-in modules expose the reference's surface."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "macbf_b200.h")


@pytest.fixture(scope="module")
def lib():
    from macbf_b200 import lib as L
    if not os.path.isfile(L.LIB_PATH):
        subprocess.check_call(["bash", os.path.join(ROOT, "macbf-gnn_b200", "build.sh")])
    return L


def header_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(macbf_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    names = header_functions()
    assert len(names) >= 15
    handle = lib.load()
    for n in names:
        assert hasattr(handle, n), f"{n} declared in the header but not exported"
    assert sorted(lib.SIGNATURES) == names, "ctypes table and header disagree"


def test_housekeeping_and_validation_without_gpu(lib):
    L = lib.load()
    assert L.macbf_version() >= 100
    assert L.macbf_neighbor_count(32, 0, 12) == 12
    assert L.macbf_neighbor_count(8, 0, 12) == 8
    assert L.macbf_neighbor_count(8, 3, 12) == 11
    # bad extents are rejected before any CUDA call, with a message
    rc = L.macbf_neighbor_topk(None, None, 0, 32, 0, 12, None, None)
    assert rc != 0 and b"B=0" in L.macbf_last_error()
    rc = L.macbf_neighbor_topk(None, None, 1, 32, 0, 129, None, None)
    assert rc != 0 and b"top_k" in L.macbf_last_error()
    rc = L.macbf_cbf_forward(None, None, None, None, 1, 32, 0, 12, 1, 0.07, 1.0, None, None)
    assert rc != 0 and b"NULL" in L.macbf_last_error()
    rc = L.macbf_cbf_forward(None, None, None, None, 1, 8, 0, 12, 1, 0.07, 1.0, None, None)
    assert rc != 0 and b"exceeds" in L.macbf_last_error()
    assert L.macbf_loss_scratch_bytes(64, 1024, 12) == 64 * 12 * 10 * 4
    assert L.macbf_step_workspace_bytes(64, 1024, 0, 12) >= 64 * 1024 * 128 * 4


def test_no_cpu_fallback(lib):
    import torch
    from macbf_b200 import ops
    with pytest.raises(lib.MacbfError):
        ops.neighbor_topk(torch.zeros(1, 32, 4), 12)


def test_dropin_surface():
    import cbf, config, controller, core
    for name in ("generate_obstacle_circle", "generate_obstacle_rectangle", "generate_data", "dynamics",
                 "loss_barrier", "loss_derivatives", "loss_actions", "ttc_dangerous_mask", "ttc_dangerous_mask_np",
                 "remove_distant_agents", "TOP_K", "TIME_STEP", "DIST_MIN_THRES"):
        assert hasattr(core, name), name
    ref = dict(TIME_STEP=0.1, WEIGHT_DECAY=1e-8, ALPHA_CBF=1.0, DIST_MIN_THRES=0.07, DIST_MIN_CHECK=0.07,
               DIST_MIN_ENLARGED=0.07, OBS_RADIUS=1.0, TOP_K=12, TIME_TO_COLLISION=2.0, TIME_TO_COLLISION_CHECK=0.1,
               TRAIN_STEPS=70000, EVALUATE_STEPS=10, INNER_LOOPS=50, REFINE_LOOPS=50, REFINE_LEARNING_RATE=0.3,
               LEARNING_RATE=1e-4, DISPLAY_STEPS=200, SAVE_STEPS=1000, ADD_NOISE_PROB=0.0, NOISE_SCALE=0.3)
    for k, v in ref.items():
        assert getattr(config, k) == v, k
    import torch
    torch.manual_seed(0)
    c, k = cbf.CBF(4), controller.Controller(4)
    z = np.load(os.path.join(ROOT, "tests", "golden", "ref_c1_n32.npz"))
    # same keys, shapes AND initial values as the reference under the same seed
    for n, v in c.state_dict().items():
        assert np.array_equal(z["cbf/" + n], v.numpy()), n
    for n, v in k.state_dict().items():
        assert np.array_equal(z["ctrl/" + n], v.numpy()), n
    assert sum(p.numel() for p in c.parameters()) == 17089
    assert sum(p.numel() for p in k.parameters()) == 34052


def test_host_generators_match_reference():
    import core
    z = np.load(os.path.join(ROOT, "tests", "golden", "ref_generate_data_n64_seed3.npz"))
    np.random.seed(3)
    s, g = core.generate_data(64, 0.07)
    assert np.array_equal(s, z["states"]) and np.array_equal(g, z["goals"])
    c = core.generate_obstacle_circle((1.0, 2.0), 0.3)
    assert c.shape == (12, 2) and np.allclose(np.linalg.norm(c - np.array([1.0, 2.0]), axis=1), 0.3)
    r = core.generate_obstacle_rectangle((0.0, 0.0), (0.6, 0.4))
    assert r.shape == (12, 2) and np.abs(r[:, 0]).max() <= 0.3 + 1e-12 and np.abs(r[:, 1]).max() <= 0.2 + 1e-12
    zz = np.load(os.path.join(ROOT, "tests", "golden", "ref_n200_vel.npz"))
    assert np.array_equal(core.ttc_dangerous_mask_np(zz["states"]), zz["ttc_np"])
    # obstacle point sets: bit-identical to the unmodified reference's (tests/golden/make_golden.py: obstacle_case)
    from helpers import OBSTACLE_SPECS
    zo = np.load(os.path.join(ROOT, "tests", "golden", "ref_obstacles.npz"))
    for i, (kind, center, size) in enumerate(OBSTACLE_SPECS):
        fn = core.generate_obstacle_circle if kind == "circle" else core.generate_obstacle_rectangle
        assert np.array_equal(fn(center, size), zo[f"p{i}"]), (kind, center, size)
    assert np.array_equal(core.generate_obstacle_circle((1.0, 2.0), 0.3, num=7), zo["circle_num7"])
    assert np.array_equal(core.generate_obstacle_rectangle((0.0, 0.0), (0.6, 0.4), num=30), zo["rect_num30"])
