"""CPU: libniti_b200.so loads without a GPU and exports every function include/niti_b200.h declares, and the ctypes
mirror (mandheling_dsp_training_b200/lib.py) binds each of them.  No compute call is made."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "niti_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)                 # comments mention function names too
    src = re.sub(r"^\s*#.*$", " ", src, flags=re.M)
    names = re.findall(r"\b(niti_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


@pytest.fixture(scope="module")
def lib():
    from mandheling_dsp_training_b200 import build
    path = build.build()                                           # cross-compiles here when stale
    return C.CDLL(path)


def test_header_declares_something():
    names = _declared()
    assert len(names) > 70 and "niti_matmul_nt_static" in names and "niti_conv_fwd_fused" in names


def test_every_declared_entry_point_is_exported(lib):
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, f"declared in include/niti_b200.h but not exported: {missing}"


def test_ctypes_mirror_covers_the_header():
    from mandheling_dsp_training_b200.lib import EXPORTED_SYMBOLS
    declared = set(_declared())
    bound = set(EXPORTED_SYMBOLS)
    assert not (bound - declared), f"bound in lib.py but not declared: {sorted(bound - declared)}"
    # device-memory, copy and layout-gather helpers are only used by the C++ plug-in; everything else must be reachable from Python
    cpp_only = {n for n in declared if n.startswith(("niti_dev_", "niti_copy_", "niti_stream_sync", "niti_gather_i8"))}
    assert not (declared - bound - cpp_only), f"declared but not bound in lib.py: {sorted(declared - bound - cpp_only)}"


def test_no_cpu_fallback_without_a_device(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    h = C.c_void_p()
    lib.niti_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    assert lib.niti_ctx_create(0, C.byref(h)) == 2 and not h.value          # NITI_ERR_CUDA, no context
