"""TEST INFRASTRUCTURE ONLY -- CPU restatement (NumPy) of xGPR's random-feature hot path.

This module is the *checker*, never the product: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import it.  Nothing under ``xgpr_b200/`` imports ``oracle``.

Parity status: PINNED.  Every function below is checked (tests/test_oracle.py)
against (a) the reference's own CPU extension compiled from /root/reference by
``oracle/Makefile`` into ``oracle/_ref/`` and (b) the golden vectors under
``tests/golden/`` that were generated from that extension by
``tests/golden/make_golden.py``.

All citations are file:line under /root/reference/src/xGPR/random_feature_generation/
(``cpu/`` = ``cpu_rf_gen/``).  ``T`` is the FHT dtype (np.float32 in production,
np.float64 for double_precision_fht); outputs are float64 (float32 for max-pool).
"""
from __future__ import annotations

import math

import numpy as np


# --------------------------------------------------------------------------- helpers
def padded_width(d: int) -> int:
    """P = 2^ceil(log2(max(d,2))).  cpu/rbf_ops/rbf_ops.cpp:56-59,
    cpu/convolution_ops/rbf_convolution.cpp:60-64."""
    return int(2 ** math.ceil(math.log2(max(int(d), 2))))


def norm_constant(p: int, dtype) -> np.generic:
    """P^-1/2 as the reference rounds it: ``T c = log2(P)/2; c = 1/pow(2,c)``.
    cpu/shared_fht_functions/shared_rfgen_ops.cpp:54-55 (also :22-23)."""
    t = np.dtype(dtype).type
    c = t(math.log2(p) / 2.0)
    return t(1.0 / math.pow(2.0, float(c)))


def fwht_(x: np.ndarray) -> np.ndarray:
    """Unnormalised in-place FWHT over the last axis, stages h = 1, 2, 4, ... P/2 in
    that order, butterfly (a, b) -> (a + b, a - b) with a at the lower index.
    cpu/shared_fht_functions/hadamard_transforms.cpp:82-127 (singleVectorTransform),
    :16-71 (transformRows)."""
    p = x.shape[-1]
    if p < 2 or (p & (p - 1)) != 0:
        raise RuntimeError("last dim not power of 2 > 1")
    lead = x.shape[:-1]
    h = 1
    while h < p:
        v = x.reshape(lead + (p // (2 * h), 2, h))
        a = v[..., 0, :].copy()
        b = v[..., 1, :]
        v[..., 0, :] = a + b
        v[..., 1, :] = a - b
        h <<= 1
    return x


def srht_(x: np.ndarray, radem: np.ndarray) -> np.ndarray:
    """x <- FWHT(x * radem * P^-1/2), rows of a 2-D array.
    cpu/basic_ops/transform_functions.cpp:94-121; shared_rfgen_ops.cpp:18-34."""
    if x.ndim != 2 or x.shape[1] != radem.shape[0]:
        raise RuntimeError("incorrect array dims passed")
    t = x.dtype.type
    c = norm_constant(x.shape[1], x.dtype)
    x *= (radem.astype(x.dtype) * c).astype(x.dtype)[None, :]
    assert x.dtype.type is t
    return fwht_(x)


def sorf_(buf: np.ndarray, radem: np.ndarray, rep: int) -> np.ndarray:
    """buf <- H D3 H D2 H D1 buf on the last axis; D_s = diag(radem[s,0,rep*P:(rep+1)*P])*c.
    cpu/shared_fht_functions/shared_rfgen_ops.cpp:50-78 (singleVectorSORF)."""
    p = buf.shape[-1]
    c = norm_constant(p, buf.dtype)
    for s in range(3):
        d = (radem[s, 0, rep * p:(rep + 1) * p].astype(buf.dtype) * c).astype(buf.dtype)
        buf *= d
        fwht_(buf)
    return buf


def _pad_rows(x2d: np.ndarray, p: int) -> np.ndarray:
    out = np.zeros((x2d.shape[0], p), dtype=x2d.dtype)
    out[:, :x2d.shape[1]] = x2d
    return out


def _check_sincos(num_rffs, num_freqs, radem, p):
    # cpu/rbf_ops/rbf_ops.cpp:49-62
    if num_rffs < 2 or (num_rffs & 1) != 0:
        raise RuntimeError("last dim of output must be even number")
    if 2 * num_freqs != num_rffs or num_freqs > radem.shape[2]:
        raise RuntimeError("incorrect number of rffs and or freqs.")
    if radem.shape[2] % p != 0:
        raise RuntimeError("incorrect number of rffs and or freqs.")


def _rbf_post(buf, chi, out, rep, scale):
    """singleVectorRBFPostProcess, shared_rfgen_ops.cpp:91-114: a = buf*chi in T,
    cos/sin in T, widened and scaled in double, accumulated."""
    p = buf.shape[-1]
    f = chi.shape[0]
    lo = rep * p
    n = min(f, lo + p) - lo
    a = buf[..., :n] * chi[lo:lo + n]
    out[..., 2 * lo:2 * (lo + n):2] += np.cos(a).astype(np.float64) * scale
    out[..., 2 * lo + 1:2 * (lo + n):2] += np.sin(a).astype(np.float64) * scale


def _rbf_post_grad(buf, chi, out, grad, sigma, rep, scale):
    """singleVectorRBFPostGrad, shared_rfgen_ops.cpp:124-156.  Note cosVal/sinVal and the
    gradient products are rounded to T before being accumulated in double."""
    t = buf.dtype.type
    p = buf.shape[-1]
    f = chi.shape[0]
    lo = rep * p
    n = min(f, lo + p) - lo
    g = buf[..., :n] * chi[lo:lo + n]                       # T
    pv = (g.astype(np.float64) * sigma).astype(t)           # T <- T*double
    cv = (np.cos(pv).astype(np.float64) * scale).astype(t)  # T <- T*double
    sv = (np.sin(pv).astype(np.float64) * scale).astype(t)
    out[..., 2 * lo:2 * (lo + n):2] += cv.astype(np.float64)
    out[..., 2 * lo + 1:2 * (lo + n):2] += sv.astype(np.float64)
    grad[..., 2 * lo:2 * (lo + n):2] -= (sv * g).astype(np.float64)
    grad[..., 2 * lo + 1:2 * (lo + n):2] += (cv * g).astype(np.float64)


# --------------------------------------------------------------------------- RBF family
def rbf_feature_gen(x, radem, chi, fit_intercept, out=None):
    """cpuRBFFeatureGen / rbfFeatureGen_<T>: cpu/rbf_ops/rbf_ops.cpp:26-106.
    x (N,D) T; radem (3,1,R) int8; chi (F) T; returns/accumulates out (N,2F) f64."""
    n, d = x.shape
    f = chi.shape[0]
    if out is None:
        out = np.zeros((n, 2 * f), np.float64)
    if n == 0 or out.shape[0] != n:
        raise RuntimeError("no datapoints")
    p = padded_width(d)
    _check_sincos(out.shape[1], f, radem, p)
    t = x.dtype.type
    # rbf_ops.cpp:64-69 -- the constant is declared T, then widened at the call site.
    scale = float(t(math.sqrt(1.0 / (f - 0.5)))) if fit_intercept else float(t(math.sqrt(1.0 / f)))
    reps = (f + p - 1) // p
    for k in range(reps):
        buf = _pad_rows(x, p)
        sorf_(buf, radem, k)
        _rbf_post(buf, chi, out, k, scale)
    return out


def rbf_grad(x, radem, chi, sigma, fit_intercept, out=None, grad=None):
    """cpuRBFGrad / rbfGrad_<T>: cpu/rbf_ops/rbf_ops.cpp:135-221.  Input is NOT pre-scaled by
    sigma; the scale constant is a genuine double here (:179)."""
    n, d = x.shape
    f = chi.shape[0]
    if out is None:
        out = np.zeros((n, 2 * f), np.float64)
    if grad is None:
        grad = np.zeros((n, 2 * f, 1), np.float64)
    if n == 0 or out.shape[0] != n:
        raise RuntimeError("no datapoints")
    p = padded_width(d)
    _check_sincos(out.shape[1], f, radem, p)
    if grad.shape[0] != out.shape[0] or grad.shape[1] != out.shape[1]:
        raise RuntimeError("Wrong array sizes.")
    scale = math.sqrt(1.0 / (f - 0.5)) if fit_intercept else math.sqrt(1.0 / f)
    reps = (f + p - 1) // p
    g2 = grad[:, :, 0]
    for k in range(reps):
        buf = _pad_rows(x, p)
        sorf_(buf, radem, k)
        _rbf_post_grad(buf, chi, out, g2, float(sigma), k, scale)
    return out, grad


# --------------------------------------------------------------------------- convolution
def _check_conv(x, seqlen, conv_width):
    # cpu/convolution_ops/rbf_convolution.cpp:55-82
    if seqlen.shape[0] != x.shape[0]:
        raise RuntimeError("wrong array sizes")
    if x.shape[1] < conv_width or conv_width <= 0:
        raise RuntimeError("invalid conv_width")
    if seqlen.max() > x.shape[1] or seqlen.min() < conv_width:
        raise RuntimeError("All sequence lengths must be >= conv width and < array size.")


def _row_scale(f, nk, scaling_type):
    # rbf_convolution.cpp:40,100-111; constants cpu/convolution_ops/rbf_convolution.h:21-23
    base = math.sqrt(1.0 / float(f))
    if scaling_type == 1:
        return base / np.sqrt(nk.astype(np.float64))
    if scaling_type == 2:
        return base / nk.astype(np.float64)
    return np.full(nk.shape, base, np.float64)


def conv_rbf_feature_gen(x, radem, chi, seqlen, conv_width, scaling_type, out=None):
    """cpuConv1dFGen / convRBFFeatureGen_<T>: cpu/convolution_ops/rbf_convolution.cpp:22-139.
    x (N,M,C) T; window j of row i is x[i, j:j+w, :] flattened, zero padded to P; sum-pooled."""
    n, m, c = x.shape
    f = chi.shape[0]
    if out is None:
        out = np.zeros((n, 2 * f), np.float64)
    if n == 0 or out.shape[0] != n:
        raise RuntimeError("no datapoints")
    p = padded_width(conv_width * c)
    _check_sincos(out.shape[1], f, radem, p)
    _check_conv(x, seqlen, conv_width)
    nk = seqlen.astype(np.int64) - conv_width + 1
    scale = _row_scale(f, nk, scaling_type)[:, None]
    reps = (f + p - 1) // p
    flat = x.reshape(n, m * c)
    for j in range(int(nk.max())):
        rows = np.nonzero(nk > j)[0]
        win = flat[rows, j * c:(j + conv_width) * c]
        sub = np.zeros((rows.size, 2 * f), np.float64)
        for k in range(reps):
            buf = _pad_rows(win, p)
            sorf_(buf, radem, k)
            _rbf_post(buf, chi, sub, k, scale[rows])
        out[rows] += sub
    return out


def conv_rbf_grad(x, radem, chi, seqlen, sigma, conv_width, scaling_type, out=None, grad=None):
    """cpuConvGrad / convRBFGrad_<T>: cpu/convolution_ops/rbf_convolution.cpp:159-282."""
    n, m, c = x.shape
    f = chi.shape[0]
    if out is None:
        out = np.zeros((n, 2 * f), np.float64)
    if grad is None:
        grad = np.zeros((n, 2 * f, 1), np.float64)
    if n == 0 or out.shape[0] != n:
        raise RuntimeError("no datapoints")
    p = padded_width(conv_width * c)
    _check_sincos(out.shape[1], f, radem, p)
    _check_conv(x, seqlen, conv_width)
    if grad.shape[0] != out.shape[0] or grad.shape[1] != out.shape[1]:
        raise RuntimeError("wrong array sizes")
    nk = seqlen.astype(np.int64) - conv_width + 1
    scale = _row_scale(f, nk, scaling_type)[:, None]
    reps = (f + p - 1) // p
    flat = x.reshape(n, m * c)
    for j in range(int(nk.max())):
        rows = np.nonzero(nk > j)[0]
        win = flat[rows, j * c:(j + conv_width) * c]
        sub = np.zeros((rows.size, 2 * f), np.float64)
        gsub = np.zeros((rows.size, 2 * f), np.float64)
        for k in range(reps):
            buf = _pad_rows(win, p)
            sorf_(buf, radem, k)
            _rbf_post_grad(buf, chi, sub, gsub, float(sigma), k, scale[rows])
        out[rows] += sub
        grad[rows, :, 0] += gsub
    return out, grad


def conv_maxpool(x, radem, chi, seqlen, conv_width, out=None):
    """cpuConv1dMaxpool / conv1dMaxpoolFeatureGen_<T>:
    cpu/convolution_ops/conv1d_operations.cpp:22-124 and :145-168.  out (N,F) float32,
    zero-initialised by the caller (=> ReLU); F == num_rffs; R == reps*P exactly."""
    n, m, c = x.shape
    f = chi.shape[0]
    if out is None:
        out = np.zeros((n, f), np.float32)
    if n == 0 or out.shape[0] != n:
        raise RuntimeError("no datapoints")
    if out.shape[1] < 2 or (out.shape[1] & 1) != 0:
        raise RuntimeError("last dim of output must be even number")
    if f != out.shape[1] or f > radem.shape[2]:
        raise RuntimeError("incorrect number of rffs and or freqs.")
    _check_conv(x, seqlen, conv_width)
    p = padded_width(conv_width * c)
    reps = (f + p - 1) // p
    if radem.shape[2] != reps * p:
        raise RuntimeError("incorrect number of rffs and or freqs.")
    nk = seqlen.astype(np.int64) - conv_width + 1
    flat = x.reshape(n, m * c)
    for j in range(int(nk.max())):
        rows = np.nonzero(nk > j)[0]
        win = flat[rows, j * c:(j + conv_width) * c]
        for k in range(reps):
            buf = _pad_rows(win, p)
            sorf_(buf, radem, k)
            lo = k * p
            cnt = min(f, lo + p) - lo
            prod = (buf[:, :cnt] * chi[lo:lo + cnt]).astype(np.float32)
            out[rows, lo:lo + cnt] = np.maximum(out[rows, lo:lo + cnt], prod)
    return out


# --------------------------------------------------------------------------- MiniARD gradient
def ard_grad(x, precomp_weights, sigma_map, sigma_vals, fit_intercept, num_lengthscales=None,
             out=None, grad=None):
    """cpuMiniARDGrad / ardGrad_<T>: cpu/rbf_ops/ard_ops.cpp:39-125.
    d_k = x_k * W[f,k] (rounded to T); G[f,l] = sum_{map[k]=l} d_k; rf = sum_k sigma_k d_k (double);
    out[2f],out[2f+1] = scale*(cos rf, sin rf); grad[2f,l] = -G*sin*scale; grad[2f+1,l] = G*cos*scale.
    scale is declared T (:89-94)."""
    n, d = x.shape
    f = precomp_weights.shape[0]
    if num_lengthscales is None:
        num_lengthscales = int(sigma_map.max()) + 1 if grad is None else grad.shape[2]
    L = num_lengthscales
    if out is None:
        out = np.zeros((n, 2 * f), np.float64)
    if grad is None:
        grad = np.zeros((n, 2 * f, L), np.float64)
    if n == 0 or out.shape[0] != n:
        raise RuntimeError("no datapoints")
    if grad.shape[0] != out.shape[0] or grad.shape[1] != out.shape[1]:
        raise RuntimeError("Wrong array sizes.")
    if precomp_weights.shape[1] != d or out.shape[1] != 2 * f or sigma_map.shape[0] != d \
            or sigma_vals.shape[0] != d:
        raise RuntimeError("Wrong array sizes.")
    t = x.dtype.type
    scale = float(t(math.sqrt(1.0 / (f - 0.5)))) if fit_intercept else float(t(math.sqrt(1.0 / f)))
    # sequential-k accumulation order, as the triple loop does it (ard_ops.cpp:98-108)
    g = grad.reshape(n, f, 2, L)
    acc = g[:, :, 0, :].copy()
    rf = np.zeros((n, f), np.float64)
    for k in range(d):
        dk = (x[:, k:k + 1] * precomp_weights[None, :, k]).astype(np.float64)   # product in T
        acc[:, :, sigma_map[k]] += dk
        rf += sigma_vals[k] * dk
    cosv = scale * np.cos(rf)
    sinv = scale * np.sin(rf)
    out[:, 0::2] = cosv
    out[:, 1::2] = sinv
    g[:, :, 0, :] = -acc * sinv[:, :, None]
    g[:, :, 1, :] = acc * cosv[:, :, None]
    return out, grad


# --------------------------------------------------------------------------- downstream algebra
def zty_ztz(z, y):
    """Z^T y, Z^T Z, y^T y as the reference accumulates them per chunk in float64.
    src/xGPR/scoring_toolkit/exact_nmll_calcs.py:68-72."""
    return z.T @ y, z.T @ z, float(y.T @ y)


def cg_matvec_chunk(z, v):
    """One chunk's contribution to the CG matvec: Z^T (Z v).
    src/xGPR/fitting_toolkit/cg_tools.py:42-44."""
    return z.T @ (z @ v)
