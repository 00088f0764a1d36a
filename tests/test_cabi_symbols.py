"""CPU-side checks of the boundary: the library loads, exports every symbol include/sb200.h declares, the Python
binding covers them, and host-only entry points / argument validation behave (no kernel is launched)."""
import ctypes
import os
import re

import numpy as np
import pytest

from seldon_core_b200 import _cabi
from seldon_core_b200.errors import SeldonB200Error

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "sb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = _cabi.lib()
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), "libsb200.so does not export %s" % n
        assert n in _cabi.SIGNATURES, "python binding misses %s" % n
    assert sorted(_cabi.SIGNATURES) == names


def test_header_cites_the_reference():
    text = open(os.path.join(ROOT, "include", "sb200.h")).read()
    for cite in ("CoreMahalanobis.py", "EpsilonGreedy.py", "ThompsonSampling.py", "utils.py:161-169",
                 "AverageCombinerUnit.java", "MnistCombiner.py"):
        assert cite in text


def test_host_only_entry_points():
    L = _cabi.lib()
    assert L.sb_version() == 100
    from oracle import codec

    for cols in (1, 4, 11, 12, 64, 256, 1000):
        assert L.sb_listvalue_nbytes(3, cols) == 3 * codec.listvalue_row_layout(cols)[0]
    assert L.sb_listvalue_nbytes(7, 0) == 77
    for shape in ((2, 3), (10000, 4), (1, 1), (16777216, 64)):
        assert L.sb_tensor_values_offset(*shape) == codec.tensor_values_offset(shape)


def test_argument_validation_without_gpu():
    L = _cabi.lib()
    h = ctypes.c_void_p()
    assert L.sb_maha_create(ctypes.byref(h), 0, 3, 25.0, 3.0, 50, -1) == -1
    assert b"outside" in L.sb_last_error()
    assert L.sb_maha_create(ctypes.byref(h), 64, 40, 25.0, 3.0, 50, -1) == -1     # > 32 components
    assert L.sb_maha_create(ctypes.byref(h), 4, 3, 25.0, 3.0, 50, 1) == -1        # max_n = 1
    assert b"ZeroDivisionError" in L.sb_last_error()
    assert L.sb_combiner_mean(None, 0, 10, None, 0) == -1
    assert b"Combiner received no inputs" in L.sb_last_error()
    r = ctypes.c_void_p()
    assert L.sb_router_create(ctypes.byref(r), 0, 0, 0.1, 1, 0) == -1
    assert L.sb_router_create(ctypes.byref(r), 0, 3, 0.1, 0, 0) == -1             # unseeded is a host decision
    with pytest.raises(SeldonB200Error):
        _cabi.check(-1)


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import seldon_core_b200 as s

    with pytest.raises(SeldonB200Error) as e:
        s.OutlierMahalanobisB200().predict(np.zeros((3, 4)), None)
    assert e.value.status_code == 500 and "no CPU fallback" in e.value.message
    with pytest.raises(SeldonB200Error):
        s.MeanCombinerB200().aggregate([np.zeros((2, 2)), np.zeros((2, 2))], [])
    with pytest.raises(SeldonB200Error):
        s.EpsilonGreedyB200(n_branches=3, seed=1).route(None, None)


def test_product_code_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "seldon_core_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                assert "libsb_oracle" not in src, f
