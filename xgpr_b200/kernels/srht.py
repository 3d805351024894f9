"""SRHTCompressor: subsampled randomised Hadamard transform of the feature axis (preconditioner).

Mirrors src/xGPR/kernels/srht_compressor.py:37-65 (draws), :70-97 (transform_x), :100-131 (device).
"""
from math import ceil

import numpy as np
import torch

from .. import ext
from .base import to_device, to_host

RADEM_VALUES = np.asarray([-1, 1], dtype=np.int8)


class SRHTCompressor:
    def __init__(self, compression_size, input_size, device="cpu", double_precision=True, random_seed=123):
        self.compression_size = compression_size
        self.input_size = input_size
        if compression_size >= input_size or compression_size <= 1:
            raise RuntimeError("The compression size should be < the number of rffs and > 1.")
        self.padded_dims = 2 ** ceil(np.log2(max(input_size, 2)))
        self.double_precision = double_precision
        rng = np.random.default_rng(random_seed)
        self.radem = rng.choice(RADEM_VALUES, size=(self.padded_dims), replace=True)
        self.col_sampler = rng.permutation(self.padded_dims)
        self.truncated_sampler = self.col_sampler[:self.compression_size]
        self.device = device

    @property
    def dtype(self):
        return torch.float64 if self.double_precision else torch.float32

    def transform_x(self, features, no_compression=False):
        if self.device != "cuda":
            raise RuntimeError("xgpr_b200 runs on device='cuda' only (no CPU fallback).")
        if len(features.shape) != 2 or features.shape[1] != self.input_size:
            raise RuntimeError("Input with unexpected size passed to a compressor module.")
        if not no_compression and isinstance(features, torch.Tensor) and features.is_cuda and features.dim() == 2 \
                and ext.srht_sample_supported(self.padded_dims, self.dtype):
            # one pass: read the chunk once, write only the sampled columns (csrc/fht_kernels.cu srht_sample_kernel)
            src = features if (features.dtype == self.dtype and features.is_contiguous()) else \
                features.to(self.dtype).contiguous()
            cols = getattr(self, "_cols_i32", None)
            if cols is None or cols.device != src.device:
                cols = self._cols_i32 = self.truncated_sampler.to(device=src.device, dtype=torch.int32).contiguous()
            out = torch.empty((src.shape[0], self.compression_size), dtype=self.dtype, device=src.device)
            return ext.srht_sample(src, self.radem, cols, out)
        if features.shape[1] < self.padded_dims:
            xfeatures = torch.zeros((features.shape[0], self.padded_dims), dtype=self.dtype, device=features.device)
            xfeatures[:, :features.shape[1]] = features
        else:
            xfeatures = features.to(self.dtype, copy=True).contiguous()
        ext.cudaSRHT(xfeatures, self.radem)
        if no_compression:
            return xfeatures[:, self.col_sampler]
        return xfeatures[:, self.truncated_sampler]

    def transform_encoded(self, c, scale, fit_intercept):
        """transform_x(Z) for Z given as its float32 encoding c (Z = double(c) * scale, column 0 = 1 with an intercept):
        the same fp64 butterflies on the same values, read at 4 bytes per feature.  None when the fused kernel does not
        cover this width (the caller then materialises Z)."""
        if self.device != "cuda" or not self.double_precision or c.shape[1] != self.input_size \
                or not ext.srht_sample_supported(self.padded_dims, torch.float64):
            return None
        cols = getattr(self, "_cols_i32", None)
        if cols is None or cols.device != c.device:
            cols = self._cols_i32 = self.truncated_sampler.to(device=c.device, dtype=torch.int32).contiguous()
        out = torch.empty((c.shape[0], self.compression_size), dtype=torch.float64, device=c.device)
        return ext.srht_sample_c32(c, self.radem, cols, out, scale, fit_intercept)

    @property
    def device(self):
        return self.device_

    @device.setter
    def device(self, value):
        if value not in ("cpu", "cuda"):
            raise RuntimeError("Unrecognized device supplied. Must be one of 'cpu', 'cuda'.")
        for nm in ("radem", "col_sampler", "truncated_sampler"):
            arr = getattr(self, nm)
            setattr(self, nm, to_device(arr) if value == "cuda" else to_host(arr))
        self.device_ = value
