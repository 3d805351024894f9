"""Test-only providers that let xgpr_b200/fitting.py run on CPU tensors: features from the NumPy oracle and
the three consumers of Z as torch fp64 matmuls.  Used by the world-size-2 gloo tests; never by the product."""
import numpy as np
import torch

from oracle import sorf_oracle as so


class OracleOps:
    """Same call signatures as xgpr_b200.ext's consumers."""

    @staticmethod
    def zty_accum(z, y, zty):
        zty += z.T @ y

    @staticmethod
    def ztz_accum(z, ztz):
        ztz += z.T @ z

    @staticmethod
    def atb_accum(a, z, acc):
        acc += a.T @ z

    @staticmethod
    def cg_matvec_accum(z, v, w, scratch=None):
        w += z.T @ (z @ v)


class OracleKernel:
    """Wraps an xgpr_b200 kernel object held on device='cpu' (host parameter draws only) and generates its
    features with the oracle.  Supports RBF-family fixed-vector kernels and Conv1d/Graph sum-pool kernels."""

    def __init__(self, kern):
        self.k = kern
        self.device = "cpu"

    def __getattr__(self, name):
        return getattr(self.k, name)

    def transform_x(self, x, sequence_length=None):
        k = self.k
        dt = np.float64 if k.double_precision else np.float32
        xin = np.ascontiguousarray(np.asarray(x), dtype=dt).copy()
        xin *= k.hyperparams[1]
        if xin.ndim == 3:
            z = so.conv_rbf_feature_gen(xin, k.radem_diag, k.chi_arr, np.asarray(sequence_length, np.int32),
                                        k.conv_width, k.scaling_type)
        else:
            z = so.rbf_feature_gen(xin, k.radem_diag, k.chi_arr, k.fit_intercept)
        if k.fit_intercept:
            z[:, 0] = 1.0
        return torch.from_numpy(z)

    def gradient_x_y(self, x, y, sequence_length=None):
        """(Z, dZ/dsigma, y) on the RAW input, as the reference's gradient entry points (kernel_baseclass.py:364-378)."""
        k = self.k
        dt = np.float64 if k.double_precision else np.float32
        xin = np.ascontiguousarray(np.asarray(x), dtype=dt)
        sigma = float(k.hyperparams[1])
        if xin.ndim == 3:
            z, dz = so.conv_rbf_grad(xin, k.radem_diag, k.chi_arr, np.asarray(sequence_length, np.int32), sigma,
                                     k.conv_width, k.scaling_type)
        else:
            z, dz = so.rbf_grad(xin, k.radem_diag, k.chi_arr, sigma, k.fit_intercept)
        if k.fit_intercept:
            z[:, 0] = 1.0
            dz[:, 0, :] = 0.0
        return torch.from_numpy(z), torch.from_numpy(dz), torch.from_numpy(np.asarray(y, np.float64))

    def transform_x_y(self, x, y, sequence_length=None, classification=False):
        return self.transform_x(x, sequence_length), torch.from_numpy(np.asarray(y, np.float64))


class OracleSRHT:
    """SRHTCompressor on CPU tensors (kernels/srht.py draws + the oracle's srht_)."""

    def __init__(self, compression_size, input_size, random_seed=123, device="cpu"):
        from xgpr_b200.kernels import SRHTCompressor
        self.c = SRHTCompressor(compression_size, input_size, device="cpu", double_precision=True,
                                random_seed=random_seed)

    def transform_x(self, features, no_compression=False):
        c = self.c
        f = features.numpy()
        xf = np.zeros((f.shape[0], c.padded_dims))
        xf[:, :f.shape[1]] = f
        so.srht_(xf, c.radem)
        cols = c.col_sampler if no_compression else c.truncated_sampler
        return torch.from_numpy(np.ascontiguousarray(xf[:, cols]))
