"""CPU: the C-ABI library loads and exports every symbol include/mimi_b200.h declares; no compute."""
import os
import re

import pytest

from mimi_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported():
    hdr = open(os.path.join(ROOT, "include", "mimi_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(mb2_[a-z0-9_A-Z]+)\s*\(", hdr)))
    assert declared == sorted(capi.SYMBOLS)
    lib = capi.load()
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.mb2_version() >= 100


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.Mb2Error, match="no usable CUDA device"):
        capi.Context(0)
