"""Fixed-vector SORF kernels: RBF, Matern, Cauchy.

Mirrors src/xGPR/kernels/basic_kernels/sorf_kernel_baseclass.py:39-84 (parameter draws),
:88-101 (device move), :105-125 (transform), :136-162 (gradient); rbf.py:21-41; matern.py:26-59;
cauchy.py:21-47.  The host draws are bit-identical to the reference's given the same seed
(checked against tests/golden/l2_*.npz).
"""
from math import ceil

import numpy as np
import torch
from scipy.stats import chi

from .. import ext
from .base import KernelBaseclass

RADEM_VALUES = np.asarray([-1, 1], dtype=np.int8)


def draw_sorf_params(padded_dims, num_freqs, random_seed, double_precision, rng=None):
    """radem (3,1,R) int8 from Generator.choice, chi (F) from scipy chi.rvs(df=padded width) -- the
    exact recipe of sorf_kernel_baseclass.py:71-84 / conv_kernel_baseclass.py:85-99."""
    if rng is None:
        rng = np.random.default_rng(random_seed)
    nblocks = ceil(num_freqs / padded_dims) if padded_dims < num_freqs else 1
    radem = rng.choice(RADEM_VALUES, size=(3, 1, nblocks * padded_dims), replace=True)
    chi_arr = chi.rvs(df=padded_dims, size=num_freqs, random_state=random_seed)
    if not double_precision:
        chi_arr = chi_arr.astype(np.float32)
    return radem, chi_arr


class SORFKernelBaseclass(KernelBaseclass):
    def __init__(self, num_rffs, xdim, sine_cosine_kernel=True, random_seed=123, double_precision=False,
                 kernel_spec_parms={}):
        super().__init__(num_rffs, xdim, sine_cosine_kernel=sine_cosine_kernel,
                         double_precision=double_precision, kernel_spec_parms=kernel_spec_parms)
        if len(xdim) != 2:
            raise ValueError("The dimensionality of the input is inappropriate for the kernel you have selected.")
        padded_dims = 2 ** ceil(np.log2(max(xdim[-1], 2)))
        self.radem_diag, self.chi_arr = draw_sorf_params(padded_dims, self.num_freqs, random_seed, double_precision)

    def kernel_specific_set_device(self, new_device):
        self._move(("radem_diag", "chi_arr"), new_device)

    def kernel_specific_set_hyperparams(self):
        return

    def kernel_specific_transform(self, input_x, sequence_length=None):
        # `input_x *= sigma`: NumPy forms float32 * float64-scalar in double and rounds once
        ext.scale_input_(input_x, float(self.hyperparams[1]))
        output_x = torch.empty((input_x.shape[0], self.num_rffs), dtype=torch.float64, device=input_x.device)
        ext.cudaRBFFeatureGen(input_x, output_x, self.radem_diag, self.chi_arr, self.fit_intercept)
        return output_x

    # ---- fused featurise -> matvec (xgpr_b200 extension of the plugin interface; Z is never materialised).
    # Used by xgpr_b200.fitting for one right-hand side; equals transform_x(x) @ v / transform_x(x).T @ t.
    supports_fused_matvec = True

    def stage_scaled(self, input_x):
        """The staged, sigma-scaled device copy that both fused passes of one chunk share."""
        xin = self._stage_input(input_x)
        if xin.shape[1] > 1024:
            return None
        return ext.scale_input_(xin, float(self.hyperparams[1]))

    def scale_staged(self, xin):
        """stage_scaled for an input that `stage_stream` already copied to the device (scaled in place)."""
        if xin.shape[1] > 1024:
            return None
        return ext.scale_input_(xin, float(self.hyperparams[1]))

    def fused_zv(self, xs, v, t):
        ext.rbf_zv(xs, self.radem_diag, self.chi_arr, self.fit_intercept, v, t)

    def fused_ztv(self, xs, t, w):
        ext.rbf_ztv(xs, self.radem_diag, self.chi_arr, self.fit_intercept, t, w)

    def fused_zzv_ok(self):
        return ext.rbf_zzv_supported(self._xdim[-1], self.num_freqs)

    def fused_zzv(self, xs, v, w):
        """w += Z^T (Z v) in one pass (t reduced on chip); only when fused_zzv_ok()."""
        ext.rbf_zzv(xs, self.radem_diag, self.chi_arr, self.fit_intercept, v, w)

    # ---- resident float32 encoding of Z for the CG matvec (xgpr_b200 extension; see csrc/resident_matvec_kernels.cu)
    def feature_scale(self):
        """The constant the feature generators multiply cos / sin by (rbf_ops.cpp:64-69), as the kernels round it."""
        f = float(self.num_freqs)
        s = np.sqrt(1.0 / (f - 0.5)) if self.fit_intercept else np.sqrt(1.0 / f)
        return float(s if self.double_precision else np.float32(s))

    def resident_ok(self):
        return (not self.double_precision) and ext.cres_zzv_supported(self.num_rffs)

    def encoding_ok(self):
        """resident_generate / feature_scale describe Z exactly (float32 FHT, <= 1024 inputs, 16-byte rows)."""
        return (not self.double_precision) and self._xdim[-1] <= 1024 and self.num_rffs % 4 == 0

    def resident_encode(self, z, out):
        """out (float32) <- the unscaled cos / sin values of the feature chunk z: lossless, since z = double(c) * scale."""
        return ext.unscale_to_f32(z, self.feature_scale(), out)

    def resident_generate(self, input_x, out):
        """out (float32, rows x num_rffs) <- the encoding of transform_x(input_x), straight from the feature kernel (no
        fp64 round trip).  Same staging and sigma scaling as transform_x."""
        xin = self._stage_input(input_x)
        ext.scale_input_(xin, float(self.hyperparams[1]))
        ext.cudaRBFFeatureGenC32(xin, out, self.radem_diag, self.chi_arr)
        return out

    def resident_zzv(self, c, v, w):
        """w += Z^T (Z v) from the resident encoding c, one pass."""
        scale = self.feature_scale()
        vs = v * scale
        if self.fit_intercept:
            vs[0] = v[0]
        ext.cres_zzv(c, vs, w, scale, self.fit_intercept)

    def resident_zzv_scaled(self, c, vs, w, skip=None):
        """resident_zzv with the direction already scaled (vs = v * feature_scale(), vs[0] = v[0] with an intercept) and an
        optional device-side gate: the form the stream-ordered CG uses (no temporaries, capturable in a CUDA graph)."""
        ext.cres_zzv(c, vs, w, self.feature_scale(), self.fit_intercept, skip)

    def kernel_specific_gradient(self, input_x, sequence_length=None):
        output_x = torch.empty((input_x.shape[0], self.num_rffs), dtype=torch.float64, device=input_x.device)
        dz_dsigma = torch.empty((input_x.shape[0], self.num_rffs, 1), dtype=torch.float64, device=input_x.device)
        ext.cudaRBFGrad(input_x, output_x, dz_dsigma, self.radem_diag, self.chi_arr,
                        float(self.hyperparams[1]), self.fit_intercept)
        return output_x, dz_dsigma


class RBF(SORFKernelBaseclass):
    """hyperparams: [lambda, sigma] (rbf.py:38-39)."""

    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False, kernel_spec_parms={}):
        super().__init__(num_rffs, xdim, sine_cosine_kernel=True, random_seed=random_seed,
                         double_precision=double_precision, kernel_spec_parms=kernel_spec_parms)
        self.hyperparams = np.ones((2))
        self.bounds = np.asarray([[1e-3, 4], [1e-6, 1e2]])
        self.device = device


def matern_rescale_(chi_arr, matern_nu, num_freqs, random_seed):
    """chi /= sqrt(chisquare(2 nu)/(2 nu)) IN PLACE on the already-cast array (matern.py:50-54)."""
    rng = np.random.default_rng(random_seed)
    chi_arr /= np.sqrt(rng.chisquare(2 * matern_nu, size=num_freqs) / (matern_nu * 2))


def cauchy_rescale_(chi_arr, num_freqs, random_seed):
    """chi *= sqrt(Exp(1)) IN PLACE (cauchy.py:39-41)."""
    rng = np.random.default_rng(random_seed)
    chi_arr *= np.sqrt(rng.exponential(size=num_freqs))


class Matern(SORFKernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False, kernel_spec_parms={}):
        super().__init__(num_rffs, xdim, sine_cosine_kernel=True, random_seed=random_seed,
                         double_precision=double_precision, kernel_spec_parms=kernel_spec_parms)
        if "matern_nu" not in kernel_spec_parms:
            raise ValueError("Tried to initialize a Matern kernel without supplying nu.")
        self.matern_nu = kernel_spec_parms["matern_nu"]
        if self.matern_nu < 1 / 2 or self.matern_nu > 5 / 2:
            raise ValueError("nu must be >= 1/2 and <= 5/2.")
        matern_rescale_(self.chi_arr, self.matern_nu, self.num_freqs, random_seed)
        self.hyperparams = np.ones((2))
        self.bounds = np.asarray([[1e-3, 1e1], [1e-6, 1e2]])
        self.device = device


class Cauchy(SORFKernelBaseclass):
    def __init__(self, xdim, num_rffs, random_seed=123, device="cpu", double_precision=False, kernel_spec_parms={}):
        super().__init__(num_rffs, xdim, sine_cosine_kernel=True, random_seed=random_seed,
                         double_precision=double_precision, kernel_spec_parms=kernel_spec_parms)
        cauchy_rescale_(self.chi_arr, self.num_freqs, random_seed)
        self.hyperparams = np.ones((2))
        self.bounds = np.asarray([[1e-3, 1e1], [1e-6, 1e2]])
        self.device = device
