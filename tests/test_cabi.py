"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/b2ntt.h
declares, validates arguments, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b2ntt.h")).read()
    return sorted(set(re.findall(r"\b(b2ntt_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(b2):
    L = b2.lib()
    names = _declared()
    assert len(names) >= 14
    for n in names:
        assert hasattr(L, n), n


def test_header_cites_reference_interfaces():
    src = open(os.path.join(ROOT, "include", "b2ntt.h")).read()
    for cite in ("additive_ntt.cuh:201-265", "nttconf.cuh:49-60", "bitslicing.cuh:89-104", "common.cu:6-43"):
        assert cite in src


def test_argument_validation_and_no_cpu_fallback(b2):
    import torch

    L = b2.lib()
    h = C.c_void_p()
    # invalid configurations are EINVAL with or without a device (nttconf.cuh:55-60)
    for args in [(32, 0, 0, 0, 1, 0), (32, 0, 29, 4, 1, 0), (64, 0, 10, 0, 1, 0), (32, 2, 10, 0, 1, 0), (32, 0, 10, 5, 1, 0),
                 (32, 0, 10, 0, 0, 0), (32, 1, 4, 0, 1, 0)]:
        assert L.b2ntt_plan_create(C.byref(h), *args) == -1, args
    assert L.b2ntt_strerror(-1) == b"invalid argument"
    if not torch.cuda.is_available():
        assert L.b2ntt_plan_create(C.byref(h), 32, 0, 10, 0, 1, 0) == -3  # B2NTT_ECUDA: nothing computes on the CPU
        assert b"cudaGetDeviceCount" in L.b2ntt_last_cuda_error()
    assert L.b2ntt_bitslice(None, 0, 48, 0, None) == -1
    assert L.b2ntt_mul(64, 0, None, None, None, 0, None) == -1
    assert L.b2ntt_mul(32, 1, None, None, None, 33, None) == -1


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "binius-ntt_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                s = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in s.lower() or f == "__init__.py" and "oracle" not in s, (f,)
