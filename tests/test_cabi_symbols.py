"""CPU: the C-ABI library loads and exports every symbol include/xgpr_b200.h declares (no compute)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "xgpr_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(xgpr_b200_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    from xgpr_b200 import _native
    names = _declared()
    assert len(names) >= 22
    lib = ctypes.CDLL(_native.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/xgpr_b200.h but not exported"
    assert sorted(_native.EXPORTED) == names
    assert lib.xgpr_b200_abi_version() == 1


def test_product_does_not_import_oracle():
    """The product package must never reach into oracle/ (or fall back to the CPU)."""
    pkg = os.path.join(ROOT, "xgpr_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "sorf_oracle" not in src, f


def test_argument_validation_without_gpu():
    """Type checks of the binding layer fire before any native call (nanobind .noconvert() behaviour)."""
    import numpy as np
    import pytest
    import torch
    from xgpr_b200 import ext
    x = torch.zeros((4, 8), dtype=torch.float32)
    with pytest.raises(TypeError):
        ext.cudaFastHadamardTransform2D(x)            # host tensor
    with pytest.raises(TypeError):
        ext.cudaFastHadamardTransform2D(np.zeros((4, 8), np.float32))
