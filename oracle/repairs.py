"""The REPAIRED semantics as a mechanical patch of the reference.

TEST INFRASTRUCTURE ONLY - used by ``tests/golden/make_golden.py`` in the build container
(``/root/reference`` does not exist on the GPU box; nothing at run time may call this).

The reference's loss functions (core.py:89-209) crash as shipped, so "what the reference
computes" is not observable for them.  To keep REPAIRED from being merely this repository's
reading of the source, the repairs are written down here as a list of exact text
substitutions ``(file, id, old, new, count)``: ``old`` must occur in the reference file exactly
``count`` times (the generator asserts it) and is replaced by ``new``; nothing else is touched.
The patched text is executed under the import stubs of ``ref_loader`` and its outputs are
stored as ``tests/golden/ref_repaired_*.npz``.  This list therefore *is* the definition of
REPAIRED, and the oracle restatement is checked against what it produces.

Only fragments of single statements appear below (the shortest text that identifies the
defect); the reference source itself is read from ``/root/reference`` at generation time.
"""
import importlib
import os
import sys
import tempfile

from . import ref_loader

# (file, repair id of SURVEY Appendix A, old, new, occurrences)
REPAIRS = [
    # F3 - N <= k returns a bare tensor although both callers unpack two values
    ("core.py", "F3", "    if n <= k:\n        return x\n", "    if n <= k:\n        return x, None\n", 1),
    # L1 - `(-1)` is an int, not a shape tuple
    ("core.py", "L1", ", (-1))", ", (-1,))", 4),
    # L2 - the (x, indices) tuple is not unpacked; tf.split(x, 5) means 5 equal parts = size 1 in torch
    ("core.py", "L2a", "s_diff = remove_distant_agents(s_diff, TOP_K)\n", "s_diff = remove_distant_agents(s_diff, TOP_K)[0]\n", 1),
    ("core.py", "L2b", "torch.split(s_diff, 5, dim=2)", "torch.split(s_diff, 1, dim=2)", 1),
    # L3 - int.type(...)
    ("core.py", "L3a", "num_dang = num_dang.type(torch.float32)", "num_dang = float(num_dang)", 2),
    ("core.py", "L3b", "num_safe = num_safe.type(torch.float32)", "num_safe = float(num_safe)", 2),
    # L4 - torch.max(t, 0) is a reduction over dim 0; tf.maximum(t, 0) was meant
    ("core.py", "L4a", "torch.max(dang_h + eps[0], 0)", "torch.clamp(dang_h + eps[0], min=0)", 1),
    ("core.py", "L4b", "torch.max(-safe_h + eps[1], 0)", "torch.clamp(-safe_h + eps[1], min=0)", 1),
    ("core.py", "L4c", "torch.max(-dang_deriv + eps[0], 0)", "torch.clamp(-dang_deriv + eps[0], min=0)", 1),
    ("core.py", "L4d", "torch.max(-safe_deriv + eps[1], 0)", "torch.clamp(-safe_deriv + eps[1], min=0)", 1),
    # L5 - surviving TensorFlow call
    ("core.py", "L5", "tf.reduce_sum(", "torch.sum(", 2),
    # L6 - sums the values instead of the boolean just computed (metric only)
    ("core.py", "L6", "acc_safe = torch.sum(safe_h.type(torch.float32))", "acc_safe = torch.sum(acc_safe.type(torch.float32))", 1),
    # L7 - torch.greater(float, int) / torch.Tensor(1.0)
    ("core.py", "L7a", "torch.where(\n        torch.greater(num_dang, 0), acc_dang, -torch.Tensor(1.0))",
     "acc_dang if num_dang > 0 else torch.tensor(-1.0)", 1),
    ("core.py", "L7b", "torch.where(\n        torch.greater(num_safe, 0), acc_safe, -torch.Tensor(1.0))",
     "acc_safe if num_safe > 0 else torch.tensor(-1.0)", 1),
    ("core.py", "L7c", "torch.where(\n        torch.greater(num_dang, 0), acc_dang_deriv, -torch.Tensor(1.0))",
     "acc_dang_deriv if num_dang > 0 else torch.tensor(-1.0)", 1),
    ("core.py", "L7d", "torch.where(\n        torch.greater(num_safe, 0), acc_safe_deriv, -torch.Tensor(1.0))",
     "acc_safe_deriv if num_safe > 0 else torch.tensor(-1.0)", 1),
    # L8 - cbf.forward takes states and forms the differences itself
    ("core.py", "L8", "h_next = cbf.forward(x_next)", "h_next = cbf.forward(states_next)", 1),
    # L9 - torch.Tensor(np, dtype=) / torch.transpose(t)
    ("core.py", "L9a", "state_gain = -torch.Tensor(\n", "state_gain = -torch.tensor(\n", 1),
    ("core.py", "L9b", "torch.transpose(state_gain)", "state_gain.T", 1),
    # F1 - channels-last -> channels-first by reshape (memory reinterpretation) instead of a transpose
    ("cbf.py", "F1", "self.cbf_net(x.reshape(-1, x.shape[2], x.shape[1]))", "self.cbf_net(x.permute(0, 2, 1))", 1),
    ("controller.py", "F1", "self.controller_centr_net(x.reshape(-1, x.shape[2], x.shape[1]))",
     "self.controller_centr_net(x.permute(0, 2, 1))", 1),
    # F2 - pools the arg-max indices instead of the values
    ("controller.py", "F2", "torch.max(x * mask, dim=2)[1]", "torch.max(x * mask, dim=2)[0]", 1),
]


def patched_sources(reference_dir=None):
    """{file: patched text}; raises if a substitution does not match the stated number of times."""
    reference_dir = reference_dir or ref_loader.REFERENCE_DIR
    out = {}
    for fname in ("core.py", "cbf.py", "controller.py", "config.py"):
        with open(os.path.join(reference_dir, fname)) as f:
            out[fname] = f.read()
    for fname, rid, old, new, count in REPAIRS:
        found = out[fname].count(old)
        if found != count:
            raise AssertionError(f"repair {rid} ({fname}): expected {count} occurrence(s) of {old!r}, found {found}")
        out[fname] = out[fname].replace(old, new)
    return out


def load_repaired():
    """Returns (core, cbf, controller, config) modules executed from the patched text (import stubs as in
    ref_loader; the patched copies live in a temporary directory, never in this repository)."""
    ref_loader.install_stubs()
    src = patched_sources()
    tmp = tempfile.mkdtemp(prefix="macbf_repaired_")
    for fname, text in src.items():
        with open(os.path.join(tmp, fname), "w") as f:
            f.write(text)
    names = ("core", "cbf", "controller", "config")
    saved = {n: sys.modules.pop(n) for n in names if n in sys.modules}
    sys.path.insert(0, tmp)
    try:
        mods = tuple(importlib.import_module(n) for n in names)
    finally:
        sys.path.remove(tmp)
        for n in names:
            sys.modules.pop(n, None)
        sys.modules.update(saved)
    return mods


def set_top_k(mods, k):
    """TOP_K reaches core / cbf / controller through `from config import *`, i.e. as a module global of each."""
    for m in mods:
        m.TOP_K = k
