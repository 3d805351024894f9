"""pytest configuration: registers the ``gpu`` marker and puts the repo root on sys.path."""
import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_files(prefix):
    return sorted(glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def load_golden(path):
    with np.load(path) as z:
        return {k: z[k] for k in z.files}


def ref_ext():
    """The reference's own CPU extension (oracle/_ref), or None when it was not built."""
    p = os.path.join(ROOT, "oracle", "_ref")
    if p not in sys.path:
        sys.path.insert(0, p)
    try:
        import xgpr_cpu_rfgen_cpp_ext as ext
        return ext
    except ImportError:
        return None


def tol_for(dtype):
    """Feature parity tolerance: rtol 1e-5 with an atol of the same relative size (features are
    O(scale) with zero crossings); double follows the reference's np.allclose
    default for T=double plus a tiny atol for sin/cos zero crossings."""
    if np.dtype(dtype) == np.float32:
        return dict(rtol=1e-5, atol=1e-5)
    return dict(rtol=1e-5, atol=1e-8)
