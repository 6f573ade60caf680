"""The C-ABI library loads without a GPU and exports every symbol include/yoda_b200.h declares."""
import ctypes
import os
import re

import pytest
from conftest import ROOT

from yoda_scheduler_b200 import _lib
from yoda_scheduler_b200._build import SO


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "yoda_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(yoda_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    names = declared_symbols()
    assert "yoda_eval_batch" in names and len(names) >= 20
    _lib.load()
    L = ctypes.CDLL(SO)
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/yoda_b200.h but not exported"
    assert set(names) == set(_lib.SYMBOLS), set(names) ^ set(_lib.SYMBOLS)


def test_struct_sizes_match_header():
    from yoda_scheduler_b200.snapshot import NODE_REC, POD_REQ, POD_RES
    assert (NODE_REC.itemsize, POD_REQ.itemsize, POD_RES.itemsize) == (416, 32, 72)
    assert ctypes.sizeof(_lib.Config) == 24 and ctypes.sizeof(_lib.Xchg) == 24 and ctypes.sizeof(_lib.Timing) == 44


def test_no_cpu_fallback():
    """Without a CUDA device the product path fails loudly instead of computing on the CPU."""
    from conftest import has_gpu
    if has_gpu():
        pytest.skip("a GPU is present")
    from yoda_scheduler_b200.api import Context, YodaError
    with pytest.raises(YodaError) as e:
        Context(16)
    assert "no CPU fallback" in str(e.value)
