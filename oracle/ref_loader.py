"""Import the unmodified reference (``/root/reference``) for golden-vector generation.

TEST INFRASTRUCTURE ONLY - used by ``tests/golden/make_golden.py`` in the build container.
``/root/reference`` does not exist on the GPU box, so nothing at run time may call this.

The reference cannot be imported as shipped: ``core.py:2`` imports tensorflow and
``cbf.py:1`` / ``controller.py:4`` import ``turtle`` (-> tkinter); both are absent here.
Two empty stand-in modules are placed in ``sys.modules`` (the reference files stay
untouched); each needs a real ``__spec__`` or torch._dynamo's ``find_spec`` probe raises.
"""
import importlib
import importlib.machinery
import os
import sys
import types

REFERENCE_DIR = os.environ.get("MACBF_REFERENCE_DIR", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_DIR, "core.py"))


def install_stubs():
    for name in ("tensorflow", "turtle"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__spec__ = importlib.machinery.ModuleSpec(name, None)
            if name == "turtle":
                m.forward = None
            sys.modules[name] = m


def load():
    """Returns (core, cbf, controller, config) reference modules."""
    if not available():
        raise RuntimeError(f"reference not present at {REFERENCE_DIR}")
    install_stubs()
    saved = {n: sys.modules.pop(n) for n in ("core", "cbf", "controller", "config") if n in sys.modules}
    sys.path.insert(0, REFERENCE_DIR)
    try:
        mods = tuple(importlib.import_module(n) for n in ("core", "cbf", "controller", "config"))
    finally:
        sys.path.remove(REFERENCE_DIR)
        for n in ("core", "cbf", "controller", "config"):
            sys.modules.pop(n, None)
        sys.modules.update(saved)
    return mods
