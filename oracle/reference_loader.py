"""TEST INFRASTRUCTURE ONLY — runs the reference's own ``app/deepdream.py`` / ``app/main.py``
UNMODIFIED, from where they lie under /root/reference, on the fake Keras of ``oracle.keras_shim``.

This only works in the build container: /root/reference does not exist on the GPU box.  There the
oracle is ``oracle.restatement`` (pinned to this module's outputs by tests/golden/*.npz).
"""
import importlib.util
import os
import sys

REFERENCE_ROOT = os.environ.get("DCV_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "app", "deepdream.py"))


def load_deepdream():
    """Import reference app/deepdream.py by path (module name ``deepdream``, as main.py:1 expects)."""
    from oracle import keras_shim
    keras_shim.install()
    if "deepdream" in sys.modules and getattr(sys.modules["deepdream"], "__dcv_ref__", False):
        return sys.modules["deepdream"]
    path = os.path.join(REFERENCE_ROOT, "app", "deepdream.py")
    spec = importlib.util.spec_from_file_location("deepdream", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["deepdream"] = mod
    sys.dont_write_bytecode, old = True, sys.dont_write_bytecode   # /root/reference is read-only
    try:
        spec.loader.exec_module(mod)
    finally:
        sys.dont_write_bytecode = old
    mod.__dcv_ref__ = True
    return mod


def load_main():
    """Import reference app/main.py by path (builds the module-global ``model`` at import, main.py:17)."""
    load_deepdream()
    if "dcv_ref_main" in sys.modules:
        return sys.modules["dcv_ref_main"]
    path = os.path.join(REFERENCE_ROOT, "app", "main.py")
    spec = importlib.util.spec_from_file_location("dcv_ref_main", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["dcv_ref_main"] = mod
    sys.dont_write_bytecode, old = True, sys.dont_write_bytecode
    try:
        spec.loader.exec_module(mod)
    finally:
        sys.dont_write_bytecode = old
    return mod
