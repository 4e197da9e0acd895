"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports exactly the
symbols include/dcv.h declares; and it refuses to run without a GPU instead of falling back."""
import ctypes
import os
import subprocess

import pytest

from deconv_api_b200 import _native, build


def test_library_builds_and_loads():
    path = build.build()
    assert os.path.exists(path)
    lib = _native.load()
    assert lib.dcv_version() >= 100


def test_header_and_library_agree():
    declared = set(_native.header_symbols())
    assert declared, "no DCV_API declarations found in include/dcv.h"
    assert declared == set(_native.SIGNATURES), (declared ^ set(_native.SIGNATURES))
    out = subprocess.check_output(["nm", "-D", "--defined-only", build.LIB], text=True)
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    missing = declared - exported
    assert not missing, "declared in dcv.h but not exported by libdcv.so: %s" % sorted(missing)
    extra = {s for s in exported if s.startswith("dcv_")} - declared
    assert not extra, "exported but not declared: %s" % sorted(extra)


def test_header_cites_reference_lines():
    txt = open(_native.HEADER).read()
    assert txt.count("deepdream.py:") >= 15


def test_library_is_sm100a_only():
    out = subprocess.check_output(["cuobjdump", "-lelf", build.LIB], text=True)
    archs = {l.split(".")[-2] for l in out.splitlines() if ".cubin" in l}
    assert archs == {"sm_100a"}, archs


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = _native.load()
    h = ctypes.c_void_p()
    rc = lib.dcv_create(0, ctypes.byref(h))
    assert rc < 0 and not h.value
    assert b"CUDA" in lib.dcv_last_error(None) or b"device" in lib.dcv_last_error(None)
    from deconv_api_b200 import deepdream
    import numpy as np
    with pytest.raises(_native.DcvError):
        deepdream.deprocess_image(np.zeros((4, 4, 3), np.float32))


def test_product_never_imports_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dirpath, _, files in os.walk(os.path.join(root, "deconv_api_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "/root/reference" not in src, f
