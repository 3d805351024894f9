"""MiniARD: RBF with one lengthscale per contiguous block of input features.

Mirrors src/xGPR/kernels/ARD_kernels/mini_ard.py:48-114 (constructor, draws), :117-138 (split-point
checks), :141-165 (device), :168-177 (hyperparameter -> per-feature weights), :180-204 (transform),
:208-258 (precompute_weights), :261-294 (gradient).  precompute_weights pushes an identity through the
SORF transform with this package's own fp64 FWHT kernel on the device (bit-identical to the
reference's CPU FWHT: same butterfly order), so the (F, D) weight matrix matches the reference's.
"""
from math import ceil

import numpy as np
import torch

from .. import ext
from .base import KernelBaseclass, to_device, to_host
from .sorf import draw_sorf_params


class MiniARD(KernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False,
                 kernel_spec_parms={}):
        super().__init__(num_rffs, xdim, sine_cosine_kernel=True, double_precision=double_precision,
                         kernel_spec_parms=kernel_spec_parms)
        if len(self._xdim) != 2:
            raise ValueError("The dimensionality of the input is inappropriate for the kernel you have selected.")
        if "split_points" not in kernel_spec_parms:
            raise ValueError("For the MiniARD kernel, 'kernel_specific_params' must contain a list "
                             "called 'split_points'.")
        if not isinstance(kernel_spec_parms["split_points"], list):
            raise ValueError("For the MiniARD kernel, 'split_points' must be a list.")
        self.split_pts = np.sort([0] + kernel_spec_parms["split_points"] + [xdim[1]])
        self.check_split_points(xdim)

        self.hyperparams = np.ones((self.split_pts.shape[0]))
        self.bounds = np.asarray([[1e-3, 1e2]] + [[1e-6, 1e2] for _ in range(self.hyperparams.shape[0] - 1)])

        self.padded_dims = 2 ** ceil(np.log2(max(xdim[-1], 2)))
        self.nblocks = ceil(self.num_freqs / self.padded_dims) if self.padded_dims < self.num_freqs else 1
        self.radem_diag, self.chi_arr = draw_sorf_params(self.padded_dims, self.num_freqs, random_seed,
                                                         double_precision)
        self.full_ard_weights = np.zeros((xdim[-1]))
        self.ard_position_key = np.zeros((xdim[-1]), dtype=np.int32)
        self.device_ = "cpu"
        self.kernel_specific_set_hyperparams()
        self.precomputed_weights = None
        self.device = device

    def check_split_points(self, xdim):
        if self.split_pts.shape[0] - 2 < 1:
            raise ValueError("There must be at least one split point to use MiniARD.")
        if self.split_pts[0] < 0:
            raise ValueError("The first split point must be > 0.")
        if self.split_pts[-1] > xdim[1]:
            raise ValueError("The last split point must be < shape[1] of the input data.")
        if np.diff(self.split_pts).min() == 0:
            raise ValueError("At least two of the split points supplied are identical.")

    def kernel_specific_set_device(self, new_device):
        self._move(("radem_diag", "chi_arr", "precomputed_weights", "full_ard_weights", "ard_position_key"),
                   new_device)

    def kernel_specific_set_hyperparams(self):
        """per-feature weights / lengthscale index from the hyperparameters (mini_ard.py:168-177)."""
        weights = np.zeros((self._xdim[-1]))
        key = np.zeros((self._xdim[-1]), dtype=np.int32)
        for i in range(1, self.split_pts.shape[0]):
            weights[self.split_pts[i - 1]:self.split_pts[i]] = self.hyperparams[i]
            key[self.split_pts[i - 1]:self.split_pts[i]] = i - 1
        if getattr(self, "device_", "cpu") == "cuda":
            self.full_ard_weights = to_device(weights)
            self.ard_position_key = to_device(key)
        else:
            self.full_ard_weights, self.ard_position_key = weights, key

    def kernel_specific_transform(self, input_x, sequence_length=None):
        # float64 weights x T input -> float64, then back to T (mini_ard.py:191-196)
        xtrans = ext.scale_input_(input_x, col_weights=self.full_ard_weights)
        output_x = torch.empty((input_x.shape[0], self.num_rffs), dtype=torch.float64, device=input_x.device)
        ext.cudaRBFFeatureGen(xtrans, output_x, self.radem_diag, self.chi_arr, self.fit_intercept)
        return output_x

    # ---- float32 encoding of Z for the multi-RHS matvec (xgpr_b200 extension, as SORFKernelBaseclass.resident_generate)
    def feature_scale(self):
        f = float(self.num_freqs)
        s = np.sqrt(1.0 / (f - 0.5)) if self.fit_intercept else np.sqrt(1.0 / f)
        return float(s if self.double_precision else np.float32(s))

    def encoding_ok(self):
        return (not self.double_precision) and self._xdim[-1] <= 1024 and self.num_rffs % 4 == 0

    def resident_generate(self, input_x, out):
        """out (float32, rows x num_rffs) <- the unscaled cos / sin values of transform_x(input_x)."""
        xin = self._stage_input(input_x)
        xtrans = ext.scale_input_(xin, col_weights=self.full_ard_weights)
        ext.cudaRBFFeatureGenC32(xtrans, out, self.radem_diag, self.chi_arr)
        return out

    def precompute_weights(self):
        """W (F, D) = rows of chi * (H D3 H D2 H D1) restricted to the first D inputs (mini_ard.py:208-258)."""
        if self.device != "cuda":
            raise RuntimeError("xgpr_b200 generates features on device='cuda' only (no CPU fallback).")
        radem = to_host(self.radem_diag)
        chi_arr = to_host(self.chi_arr)
        p = self.padded_dims
        norm_constant = 1.0 / (2.0 ** (np.log2(p) / 2.0))
        padded_chi = np.zeros((self.nblocks * p))
        padded_chi[:chi_arr.shape[0]] = chi_arr
        blocks = []
        for i in range(self.nblocks):
            lo, hi = i * p, (i + 1) * p
            mat = np.eye(p)
            for s in range(3):
                mat *= radem[s:s + 1, 0, lo:hi] * norm_constant
                dmat = to_device(mat)
                ext.cudaFastHadamardTransform2D(dmat)
                mat = to_host(dmat)
            mat *= padded_chi[lo:hi]
            blocks.append(mat.T[:, :self._xdim[-1]])
        weights = np.vstack(blocks)[:self.num_freqs, :]
        if not self.double_precision:
            weights = weights.astype(np.float32)
        self.precomputed_weights = to_device(np.ascontiguousarray(weights))

    def kernel_specific_gradient(self, input_x, sequence_length=None):
        if self.precomputed_weights is None:
            self.precompute_weights()
        nscales = int(self.split_pts.shape[0] - 1)
        # overwrite form: no zero-fill and no read-back of the 8 M L bytes per row the reference's `+=` contract costs
        xtrans = torch.empty((input_x.shape[0], self.num_rffs), dtype=torch.float64, device=input_x.device)
        dz_dsigma = torch.empty((input_x.shape[0], self.num_rffs, nscales), dtype=torch.float64,
                                device=input_x.device)
        ext.cudaMiniARDGrad(input_x, xtrans, self.precomputed_weights, self.ard_position_key,
                            self.full_ard_weights, dz_dsigma, self.fit_intercept, overwrite=True)
        return xtrans, dz_dsigma
