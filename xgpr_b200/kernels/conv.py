"""Sliding-window (Conv1d*) and graph (Graph*) SORF kernels, sum-pooled in-kernel.

Mirrors src/xGPR/kernels/convolution_kernels/conv_kernel_baseclass.py:44-99 (constructor, averaging ->
scaling_type, parameter draws), :103-110 (device), :114-145 (transform), :155-190 (gradient);
conv1d_rbf.py:44-56, conv1d_matern.py:46-66, conv1d_cauchy.py:40-58, graph_rbf.py:20-47,
graph_matern.py:36-59, graph_cauchy.py:32-51.  Graph kernels are conv kernels with conv_width = 1.
"""
from math import ceil

import numpy as np
import torch

from .. import ext
from .base import KernelBaseclass
from .sorf import cauchy_rescale_, draw_sorf_params, matern_rescale_

_AVERAGING = {"none": 0, "sqrt": 1, "full": 2}


class ConvKernelBaseclass(KernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, double_precision=False, conv_width=9,
                 kernel_spec_parms={}):
        super().__init__(num_rffs, xdim, sine_cosine_kernel=True, double_precision=double_precision,
                         kernel_spec_parms=kernel_spec_parms)
        if len(xdim) != 3:
            raise RuntimeError("Tried to initialize a Conv1d kernel with a 2d x-array! x should be a 3d "
                               "array for Conv1d.")
        self.scaling_type = 0
        if "averaging" in kernel_spec_parms:
            if kernel_spec_parms["averaging"] not in _AVERAGING:
                raise RuntimeError("Unrecognized value for 'averaging' supplied, should be one of "
                                   "'none', 'sqrt', 'full'.")
            self.scaling_type = _AVERAGING[kernel_spec_parms["averaging"]]
        self.conv_width = conv_width
        padded_dims = 2 ** ceil(np.log2(max(conv_width * xdim[2], 2)))
        rng = np.random.default_rng(random_seed)
        # conv draws ceil(F/P)*P signs (conv_kernel_baseclass.py:90-95): same as the fixed-vector recipe
        # except when F < P, where ceil() still gives one block -- identical.
        self.radem_diag, self.chi_arr = draw_sorf_params(padded_dims, self.num_freqs, random_seed,
                                                         double_precision, rng=rng)

    def kernel_specific_set_device(self, new_device):
        self._move(("radem_diag", "chi_arr"), new_device)

    def kernel_specific_set_hyperparams(self):
        return

    def _check(self, input_x, sequence_length):
        if sequence_length is None:
            raise RuntimeError("sequence_length is required for convolution kernels.")
        if input_x.shape[2] != self._xdim[2]:
            raise RuntimeError("Unexpected input shape supplied.")

    def kernel_specific_transform(self, input_x, sequence_length):
        self._check(input_x, sequence_length)
        ext.scale_input_(input_x, float(self.hyperparams[1]))   # NumPy rounding of `input_x *= sigma`
        xtrans = torch.zeros((input_x.shape[0], self.num_rffs), dtype=torch.float64, device=input_x.device)
        ext.cudaConv1dFGen(input_x, xtrans, self.radem_diag, self.chi_arr, sequence_length,
                           self.conv_width, self.scaling_type)
        return xtrans

    def kernel_specific_gradient(self, input_x, sequence_length):
        self._check(input_x, sequence_length)
        xtrans = torch.zeros((input_x.shape[0], self.num_rffs), dtype=torch.float64, device=input_x.device)
        dz_dsigma = torch.zeros((input_x.shape[0], self.num_rffs, 1), dtype=torch.float64, device=input_x.device)
        ext.cudaConvGrad(input_x, xtrans, self.radem_diag, self.chi_arr, sequence_length, dz_dsigma,
                         float(self.hyperparams[1]), self.conv_width, self.scaling_type)
        return xtrans, dz_dsigma


def _need_width(parms):
    if "conv_width" not in parms:
        raise ValueError("conv_width must be included as a kernel-specific parameter if using a sequence kernel.")
    return parms["conv_width"]


def _need_nu(parms):
    if "matern_nu" not in parms:
        raise ValueError("Tried to initialize a Matern kernel without supplying nu.")
    nu = parms["matern_nu"]
    if nu < 1 / 2 or nu > 5 / 2:
        raise ValueError("nu must be >= 1/2 and <= 5/2.")
    return nu


class Conv1dRBF(ConvKernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False, kernel_spec_parms={}):
        super().__init__(xdim, num_rffs, random_seed, double_precision, _need_width(kernel_spec_parms),
                         kernel_spec_parms)
        self.hyperparams = np.ones((2))
        self.bounds = np.asarray([[1e-3, 5], [1e-6, 1e2]])
        self.device = device


class Conv1dMatern(ConvKernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False, kernel_spec_parms={}):
        width = _need_width(kernel_spec_parms)
        nu = _need_nu(kernel_spec_parms)
        super().__init__(xdim, num_rffs, random_seed, double_precision, width, kernel_spec_parms)
        self.matern_nu = nu
        matern_rescale_(self.chi_arr, nu, self.num_freqs, random_seed)
        self.hyperparams = np.ones((2))
        self.bounds = np.asarray([[1e-3, 5], [1e-6, 1e2]])
        self.device = device


class Conv1dCauchy(ConvKernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False, kernel_spec_parms={}):
        super().__init__(xdim, num_rffs, random_seed, double_precision, _need_width(kernel_spec_parms),
                         kernel_spec_parms)
        cauchy_rescale_(self.chi_arr, self.num_freqs, random_seed)
        self.hyperparams = np.ones((2))
        self.bounds = np.asarray([[1e-3, 1e1], [1e-6, 1e2]])
        self.device = device


class GraphRBF(ConvKernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False, kernel_spec_parms={}):
        super().__init__(xdim, num_rffs, random_seed, double_precision, 1, kernel_spec_parms)
        self.hyperparams = np.ones((2))
        self.bounds = np.asarray([[1e-3, 1e2], [1e-2, 1e2]])
        self.device = device


class GraphMatern(ConvKernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False, kernel_spec_parms={}):
        nu = _need_nu(kernel_spec_parms)
        super().__init__(xdim, num_rffs, random_seed, double_precision, 1, kernel_spec_parms)
        self.matern_nu = nu
        matern_rescale_(self.chi_arr, nu, self.num_freqs, random_seed)
        self.hyperparams = np.ones((2))
        self.bounds = np.asarray([[1e-3, 1e1], [1e-6, 1e2]])
        self.device = device


class GraphCauchy(ConvKernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False, kernel_spec_parms={}):
        super().__init__(xdim, num_rffs, random_seed, double_precision, 1, kernel_spec_parms)
        cauchy_rescale_(self.chi_arr, self.num_freqs, random_seed)
        self.hyperparams = np.ones((2))
        self.bounds = np.asarray([[1e-3, 1e1], [1e-6, 1e2]])
        self.device = device
