"""CPU-only: the C-ABI library loads and exports every symbol that include/*.h declares (no compute without a GPU)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
SO = os.path.join(ROOT, "pipeinfer_b200", "libpipeinfer_b200.so")


def declared_symbols():
    names = set()
    inc = os.path.join(ROOT, "include")
    for f in sorted(os.listdir(inc)):
        if not f.endswith(".h"):
            continue
        src = open(os.path.join(inc, f)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        src = re.sub(r"//[^\n]*", "", src)
        for m in re.finditer(r"\b((?:pib|llama|ggml_mpi)_[a-z0-9_]+)\s*\(", src):
            names.add(m.group(1))
    return sorted(names)


@pytest.fixture(scope="module")
def so():
    if not os.path.exists(SO):
        import __graft_entry__
        __graft_entry__.build()
    return C.CDLL(SO)


def test_library_exports_every_declared_symbol(so):
    names = declared_symbols()
    assert len(names) >= 15
    missing = [n for n in names if not hasattr(so, n)]
    assert not missing, missing


def test_no_cpu_fallback_without_a_device(so):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    assert so.pib_init(0) < 0
    so.pib_last_error.restype = C.c_char_p
    assert b"no CUDA device" in so.pib_last_error()
