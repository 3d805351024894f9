"""xgpr_b200 -- B200-native (sm_100a) implementation of xGPR's random-feature hot path.

Importing this package loads the native library; there is no CPU or PyTorch fallback.
"""
from . import _native  # noqa: F401  (raises ImportError when libxgpr_b200.so has not been built)
from . import ext  # noqa: F401

__version__ = "0.1.0"
