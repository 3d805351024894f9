"""Conv1dTwoLayer: ReLU/max-pooled SORF convolution feeding an RBF layer; and the stand-alone
max-pool feature extractor used by the FastConv1d static layer.

Mirrors src/xGPR/kernels/convolution_kernels/l2_conv1d.py:55-130 (constructor: radem1 then radem2 from the
SAME generator, chi1 / chi2 from separate chi.rvs streams), :133-146 (device), :156-184 (transform),
:187-221 (gradient); conv_feature_extractor.py (FHTMaxpoolConv1dFeatureExtractor).
"""
from math import ceil

import numpy as np
import torch
from scipy.stats import chi

from .. import ext
from .base import KernelBaseclass, to_device, to_host

RADEM_VALUES = np.asarray([-1, 1], dtype=np.int8)


class Conv1dTwoLayer(KernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False,
                 kernel_spec_parms={}):
        if "conv_width" not in kernel_spec_parms:
            raise ValueError("conv_width must be included as a kernel-specific parameter if using a "
                             "sequence kernel.")
        if "init_rffs" not in kernel_spec_parms:
            raise ValueError("init_rffs must be included as a kernel-specific parameter if using the "
                             "2 layer conv1d kernel.")
        if len(xdim) != 3:
            raise RuntimeError("Tried to initialize a Conv1d kernel with a 2d x-array! x should be a 3d "
                               "array for Conv1d.")
        self.init_rffs = kernel_spec_parms["init_rffs"]
        if self.init_rffs % 2 != 0:
            raise RuntimeError("Number of init rffs should be an even number.")
        super().__init__(num_rffs, xdim, sine_cosine_kernel=True, double_precision=double_precision,
                         kernel_spec_parms=kernel_spec_parms)
        self.hyperparams = np.ones((2))
        self.bounds = np.asarray([[1e-3, 5], [1e-6, 1e2]])
        rng = np.random.default_rng(random_seed)
        self.conv_width = kernel_spec_parms["conv_width"]

        p1 = 2 ** ceil(np.log2(max(self.conv_width * xdim[2], 2)))
        self.radem_diag1 = rng.choice(RADEM_VALUES, size=(3, 1, ceil(self.init_rffs / p1) * p1), replace=True)
        self.chi_arr1 = chi.rvs(df=p1, size=self.init_rffs, random_state=random_seed)

        p2 = 2 ** ceil(np.log2(max(self.init_rffs, 2)))
        nblocks = ceil(self.num_freqs / p2) if p2 < self.num_freqs else 1
        self.radem_diag2 = rng.choice(RADEM_VALUES, size=(3, 1, nblocks * p2), replace=True)
        self.chi_arr2 = chi.rvs(df=p2, size=self.num_freqs, random_state=random_seed)
        if not self.double_precision:
            self.chi_arr1 = self.chi_arr1.astype(np.float32)
            self.chi_arr2 = self.chi_arr2.astype(np.float32)
        self.device = device

    def kernel_specific_set_device(self, new_device):
        self._move(("radem_diag1", "chi_arr1", "radem_diag2", "chi_arr2"), new_device)

    def kernel_specific_set_hyperparams(self):
        return

    def _first_layer(self, input_x, sequence_length):
        if sequence_length is None:
            raise ValueError("sequence_length is required for convolution kernels.")
        if input_x.shape[2] != self._xdim[2]:
            raise RuntimeError("Unexpected input shape supplied.")
        feat = torch.zeros((input_x.shape[0], self.init_rffs), dtype=torch.float32, device=input_x.device)
        ext.cudaConv1dMaxpool(input_x, feat, self.radem_diag1, self.chi_arr1, sequence_length, self.conv_width)
        return feat

    def kernel_specific_transform(self, input_x, sequence_length):
        feat = self._first_layer(input_x, sequence_length)
        xtrans = torch.empty((feat.shape[0], self.num_rffs), dtype=torch.float64, device=feat.device)
        ext.scale_input_(feat, float(self.hyperparams[1]))      # NumPy rounding of `featurized_x *= sigma`
        ext.cudaRBFFeatureGen(feat, xtrans, self.radem_diag2, self.chi_arr2, self.fit_intercept)
        return xtrans

    def kernel_specific_gradient(self, input_x, sequence_length=None):
        feat = self._first_layer(input_x, sequence_length)
        output_x = torch.empty((feat.shape[0], self.num_rffs), dtype=torch.float64, device=feat.device)
        dz_dsigma = torch.empty((feat.shape[0], self.num_rffs, 1), dtype=torch.float64, device=feat.device)
        ext.cudaRBFGrad(feat, output_x, dz_dsigma, self.radem_diag2, self.chi_arr2,
                        float(self.hyperparams[1]), self.fit_intercept)
        return output_x, dz_dsigma
