"""Shared helpers for the parity tests."""
import glob
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# north_star tolerance: h, u and the losses within 1e-5 relative (fp32); "relative" is taken
# against the largest magnitude of the reference tensor (scale-relative), the usual reading for
# fp32 MLP outputs that cross zero.
RTOL = 1e-5


def load_npz(name):
    z = np.load(os.path.join(GOLDEN, name))
    return {k: z[k] for k in z.files}


def oracle_cases():
    return sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLDEN, "oracle_*.npz")))


def ref_cases():
    return ["ref_c1_n32.npz", "ref_n200_vel.npz", "ref_n13_vel.npz"]


def rel_err(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    scale = max(float(np.abs(b).max()), 1e-30)
    return float(np.abs(a - b).max()) / scale


def assert_close(a, b, rtol=RTOL, what=""):
    e = rel_err(a, b)
    assert e <= rtol, f"{what}: scale-relative error {e:.3e} > {rtol:.1e}"


def dev(x, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(x))
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda()


def flat_params_from_ref(z, prefix, keys):
    return np.concatenate([z[prefix + k].reshape(-1) for k in keys]).astype(np.float32)


def oracle_params(z):
    """(p_cbf, p_ctrl) dicts from the flat parameter vectors of an oracle_*.npz fixture."""
    from oracle import macbf_oracle as O

    def split(flat, shapes, keys):
        parts = torch.split(torch.from_numpy(np.asarray(flat)), [int(np.prod(sh)) for sh in shapes])
        return dict(zip(keys, [t.reshape(shp) for t, shp in zip(parts, shapes)]))
    return split(z["cbf_flat"], O.CBF_SHAPES, O.CBF_KEYS), split(z["ctrl_flat"], O.CTRL_SHAPES, O.CTRL_KEYS)


def oracle_step_on_device_graph(p_cbf, p_ctrl, s, g, obs, k, idx_next_dev, with_grad=True, min_same=0.99):
    """The oracle's full step for a comparison that never has to be skipped.

    S' is computed from u, and the device's u is ~1 ulp away from the oracle's, so a row of the S' graph whose
    K-th and (K+1)-th keys are within a few ulp can select another neighbour on the device.  Such rows are first
    PROVEN to be near-ties (the device's neighbour set and the oracle's have the same sorted keys on the oracle's S'
    within 1e-6), then the oracle evaluates h' on the device's S' neighbour sets (checker hook `indices_next`), so
    losses and gradients are compared on the same graph - always, not only when no row flipped.
    Returns (oracle outputs, fraction of identical rows)."""
    from oracle import macbf_oracle as O
    ref = O.full_step(p_cbf, p_ctrl, s, g, obs, k, O.REPAIRED, with_grad=False)
    idx_ref = O.neighbor_indices(ref["s_next"], k, obs, row_chunk=512).numpy()
    idx_dev = np.asarray(idx_next_dev).astype(np.int64)
    same = (idx_dev == idx_ref).all(axis=1)
    assert same.mean() >= min_same, f"S' graph: only {same.mean():.4f} of the rows agree"
    if not same.all():
        cand = O.candidate_states(ref["s_next"], obs)
        rows = np.nonzero(~same)[0]
        d = ref["s_next"][rows, None, :2] - cand[None, :, :2]
        keys = O.knn_keys(d)
        kd = torch.gather(keys, 1, torch.from_numpy(idx_dev[rows])).sort(dim=1)[0]
        kr = torch.gather(keys, 1, torch.from_numpy(idx_ref[rows])).sort(dim=1)[0]
        assert float((kd - kr).abs().max()) <= 1e-6, "a row of the S' graph differs and is NOT a near-tie"
    out = O.full_step(p_cbf, p_ctrl, s, g, obs, k, O.REPAIRED, with_grad=with_grad,
                      indices_next=torch.from_numpy(idx_dev))
    return out, float(same.mean())


def assert_scalars_close(sc, ref_scalars, h_scale, n_dang, n_safe, where=""):
    """The 9 loss / accuracy scalars (include/macbf_b200.h layout) against the oracle's.

    The hinge losses (0, 1, 4, 5) are means of max(+-h + eps, 0) resp. of max(-(h' - h + dt alpha h) + eps, 0): their
    absolute error is the absolute error of h, so the 1e-5 is taken relative to the magnitude of h (`h_scale` =
    max|h|), exactly as for h itself - a small mean of differences of O(1) values cannot be more accurate than the
    values.  The accuracies (2, 3, 6, 7) are counts / n: an element within rounding of zero may fall on the other side,
    so up to 3 such elements of the smaller set are allowed.  The action loss (8) is a plain mean: relative 2e-5."""
    sc = np.asarray(sc, np.float64)
    ref = np.asarray(ref_scalars, np.float64)
    for q in (0, 1, 4, 5):
        assert abs(sc[q] - ref[q]) <= 2 * RTOL * max(abs(ref[q]), float(h_scale)), (where, q, sc[q], ref[q])
    for q, n in ((2, n_dang), (3, n_safe), (6, n_dang), (7, n_safe)):
        assert abs(sc[q] - ref[q]) <= 3.0 / max(float(n), 1.0) + 2 * RTOL, (where, q, sc[q], ref[q])
    assert abs(sc[8] - ref[8]) <= 2 * RTOL * max(abs(ref[8]), 1e-3), (where, 8, sc[8], ref[8])


# obstacle specs of tests/golden/ref_obstacles.npz (made by make_golden.py: obstacle_case from the unmodified reference)
OBSTACLE_SPECS = [("circle", (1.0, 2.0), 0.3), ("rectangle", (0.0, 0.0), (0.6, 0.4)), ("circle", (-0.37, 5.11), 0.125),
                  ("rectangle", (2.5, -1.25), (0.3, 0.9)), ("rectangle", (1.0, 1.0), (1.0, 0.1)), ("circle", (0.0, 0.0), 1.0)]
