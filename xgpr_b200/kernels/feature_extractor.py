"""FHTMaxpoolConv1dFeatureExtractor: the ReLU / global max-pool convolution feature extractor (run once on the data; its
output feeds another kernel -- the first layer of a two-layer GP, static_layers/fast_conv.py:62).

Mirrors src/xGPR/kernels/convolution_kernels/conv_feature_extractor.py:19-118: constructor (seqwidth, num_rffs,
random_seed, device, conv_width), the host draws (:58-71: radem (3, 1, ceil(num_rffs / P) P) int8 from Generator.choice, chi
(num_rffs) float32 from scipy chi.rvs(df = P)), `kernel_specific_set_device` and `transform_x(input_x, sequence_length)`
-> (n, num_rffs) float32.  Features come from the in-kernel max-pool of csrc/conv_kernels.cu (CONV_MAXPOOL, bit-exact
against the reference's cpuConv1dMaxpool).  device="cuda" only generates features: there is no CPU fallback.
"""
from math import ceil

import numpy as np
import torch
from scipy.stats import chi

from .. import ext
from .base import to_device, to_host


class FHTMaxpoolConv1dFeatureExtractor:
    def __init__(self, seqwidth, num_rffs, random_seed=123, device="cpu", conv_width=9):
        rng = np.random.default_rng(random_seed)
        self.conv_width = conv_width
        self.num_rffs = num_rffs
        self.seqwidth = seqwidth
        padded_dims = 2 ** ceil(np.log2(max(conv_width * seqwidth, 2)))
        init_calc_featsize = ceil(self.num_rffs / padded_dims) * padded_dims
        self.radem_diag = rng.choice(np.asarray([-1, 1], dtype=np.int8), size=(3, 1, init_calc_featsize), replace=True)
        self.chi_arr = chi.rvs(df=padded_dims, size=self.num_rffs, random_state=random_seed).astype(np.float32)
        self.device = device

    def kernel_specific_set_device(self, new_device):
        for nm in ("radem_diag", "chi_arr"):
            arr = getattr(self, nm)
            setattr(self, nm, to_device(arr) if new_device == "cuda" else to_host(arr))

    def transform_x(self, input_x, sequence_length):
        if sequence_length is None:
            raise ValueError("sequence_length is required for convolution kernels.")
        if input_x.shape[2] != self.seqwidth:
            raise ValueError("Unexpected number of features per timepoint / sequence element on this input.")
        if self.device != "cuda":
            raise RuntimeError("xgpr_b200 generates features on device='cuda' only (no CPU fallback).")
        if isinstance(sequence_length, torch.Tensor):
            sequence_length = sequence_length.cpu().numpy()
        sequence_length = np.ascontiguousarray(sequence_length.astype(np.int32, copy=False))
        x_in = to_device(input_x, torch.float32).contiguous()
        output_x = torch.zeros((x_in.shape[0], self.num_rffs), dtype=torch.float32, device=x_in.device)
        ext.cudaConv1dMaxpool(x_in, output_x, self.radem_diag, self.chi_arr, sequence_length, self.conv_width)
        return output_x

    @property
    def device(self):
        return self.device_

    @device.setter
    def device(self, value):
        if value not in ("cpu", "cuda"):
            raise RuntimeError("Unrecognized device supplied. Must be one of 'cpu', 'cuda'.")
        self.device_ = value
        self.kernel_specific_set_device(value)
