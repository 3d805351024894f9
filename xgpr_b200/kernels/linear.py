"""Linear kernel: features are the inputs (plus an intercept column).
Mirrors src/xGPR/kernels/basic_kernels/linear.py:24-110."""
import numpy as np
import torch

from .base import KernelBaseclass


class Linear(KernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False,
                 kernel_spec_parms={}):
        icpt = kernel_spec_parms.get("intercept", True) is not False
        super().__init__(xdim[1] + 1 if icpt else xdim[1], xdim, kernel_spec_parms=kernel_spec_parms)
        if len(xdim) > 2:
            raise ValueError("The Linear kernel is only applicable for fixed vector input.")
        self.hyperparams = np.ones((1))
        self.bounds = np.asarray([[1e-3, 1e1]])
        self.device = device

    def kernel_specific_set_device(self, new_device):
        return

    def kernel_specific_set_hyperparams(self):
        return

    def kernel_specific_transform(self, input_x, sequence_length=None):
        if self.fit_intercept:
            xtrans = torch.zeros((input_x.shape[0], input_x.shape[1] + 1), dtype=torch.float64,
                                 device=input_x.device)
            xtrans[:, 1:] = input_x
            return xtrans
        return input_x.to(torch.float64)

    def kernel_specific_gradient(self, input_x, sequence_length=None):
        xtrans = self.kernel_specific_transform(input_x)
        return xtrans, torch.zeros((xtrans.shape[0], 0, 0), dtype=torch.float64, device=xtrans.device)
