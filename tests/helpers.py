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
