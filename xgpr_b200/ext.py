"""Python face of the C ABI: the eight callables of the reference's `xgpr_cuda_rfgen_cpp_ext`
(gpu_rf_gen/xgpr_cuda_rfgen_cpp_ext.cpp:20-93) with the same names, keyword names and argument
order, taking torch CUDA tensors instead of CuPy arrays.

Like nanobind's ``.noconvert()`` ndarray arguments, a dtype / rank / device / contiguity mismatch
raises TypeError before the native call; shape violations raise RuntimeError from the native side.
Kernels are launched on ``torch.cuda.current_stream()``; nothing synchronises.
"""
import numpy as np
import torch

from . import _native
from ._native import lib, check

_SFX = {torch.float32: "_f32", torch.float64: "_f64"}


def _env_level():
    v = __import__("os").environ.get("XGPR_B200_ACCURATE_SINCOS", "")
    try:
        return int(v) if v != "" else -1
    except ValueError:
        return -1


_ACCURATE = {"level": _env_level(), "devices": set()}


def set_accurate_sincos(level):
    """float32 cos / sin of the feature kernels (include/xgpr_b200.h): 0 / False = Cody-Waite + MUFU (<= ~4e-7 absolute),
    1 / True = polynomial on the FMA pipe (<= 7e-8), 2 = CUDA's sincosf, -1 / None = the default (0).  Also selectable with XGPR_B200_ACCURATE_SINCOS=<level>.  The setting
    is a per-device constant: it is (re)applied lazily to every device the next call launches on."""
    _ACCURATE["level"] = -1 if level is None else int(level)
    _ACCURATE["devices"].clear()


def _stream():
    dev = torch.cuda.current_device()
    if dev not in _ACCURATE["devices"]:
        check(lib.xgpr_b200_set_accurate_sincos(int(_ACCURATE["level"])))
        _ACCURATE["devices"].add(dev)
    return torch.cuda.current_stream().cuda_stream


def _dev(t, name, dtypes, ndim):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise TypeError(f"{name}: expected a CUDA torch.Tensor")
    if t.dtype not in dtypes:
        raise TypeError(f"{name}: dtype {t.dtype} not accepted (expected one of {dtypes})")
    if t.dim() != ndim:
        raise TypeError(f"{name}: expected {ndim} dims, got {t.dim()}")
    if not t.is_contiguous():
        raise TypeError(f"{name}: must be C-contiguous")
    return t


def _radem3(t):
    _dev(t, "radem", (torch.int8,), 3)
    if t.shape[0] != 3 or t.shape[1] != 1:
        raise TypeError("radem: expected shape (3, 1, R)")
    return t


def _host_i32(a, name):
    if isinstance(a, torch.Tensor):
        if a.is_cuda:
            raise TypeError(f"{name}: must live on the host (as in the reference API)")
        a = a.numpy()
    if not isinstance(a, np.ndarray) or a.dtype != np.int32 or a.ndim != 1 or not a.flags["C_CONTIGUOUS"]:
        raise TypeError(f"{name}: expected a C-contiguous 1-d host int32 array")
    return a


def _same(a, b, an, bn):
    if a.dtype != b.dtype:
        raise TypeError(f"{an} and {bn} must share a dtype")


def cudaFastHadamardTransform2D(inputArr):
    x = _dev(inputArr, "inputArr", _SFX, 2)
    check(getattr(lib, "xgpr_b200_fht2d" + _SFX[x.dtype])(x.data_ptr(), x.shape[0], x.shape[1], _stream()))
    return 0


def cudaSRHT(inputArr, radem):
    x = _dev(inputArr, "inputArr", _SFX, 2)
    r = _dev(radem, "radem", (torch.int8,), 1)
    check(getattr(lib, "xgpr_b200_srht" + _SFX[x.dtype])(x.data_ptr(), x.shape[0], x.shape[1], r.data_ptr(),
                                                         r.shape[0], _stream()))
    return 0


def srht_sample_supported(p_full, dtype):
    return bool(lib.xgpr_b200_srht_sample_supported(int(p_full), 8 if dtype == torch.float64 else 4))


def srht_sample(z, radem, cols, out):
    """out (n, rank) = SRHT(z zero-padded to len(radem) columns)[:, cols] in one pass over z (z is not modified):
    SRHTCompressor.transform_x fused (srht_compressor.py:70-97).  cols: int32 device vector."""
    z = _dev(z, "z", _SFX, 2)
    r = _dev(radem, "radem", (torch.int8,), 1)
    c = _dev(cols, "cols", (torch.int32,), 1)
    o = _dev(out, "out", (z.dtype,), 2)
    if o.shape[0] != z.shape[0] or o.shape[1] != c.shape[0] or z.shape[1] > r.shape[0]:
        raise RuntimeError("wrong array sizes")
    check(getattr(lib, "xgpr_b200_srht_sample" + _SFX[z.dtype])(z.data_ptr(), z.shape[0], z.shape[1], r.data_ptr(),
                                                               r.shape[0], c.data_ptr(), c.shape[0], o.data_ptr(),
                                                               _stream()))
    return out


def scale_input_(x, sigma=1.0, col_weights=None):
    """In place x <- T(double(x) * sigma) (or * col_weights[j]): the reference's host-side
    `input_x *= hyperparams[1]` / MiniARD `x * full_ard_weights` with NumPy's rounding (one rounding of
    the double product), see include/xgpr_b200.h."""
    if x.dtype not in _SFX or not x.is_cuda or not x.is_contiguous():
        raise TypeError("x: expected a C-contiguous CUDA float32/float64 tensor")
    wptr = 0
    if col_weights is not None:
        w = _dev(col_weights, "col_weights", (torch.float64,), 1)
        if w.shape[0] != x.shape[-1]:
            raise RuntimeError("wrong array sizes")
        wptr = w.data_ptr()
    d = x.shape[-1]
    check(getattr(lib, "xgpr_b200_scale_input" + _SFX[x.dtype])(x.data_ptr(), x.numel() // d, d, wptr,
                                                                float(sigma), _stream()))
    return x


def cudaRBFFeatureGen(inputArr, outputArr, radem, chiArr, fitIntercept):
    x = _dev(inputArr, "inputArr", _SFX, 2)
    o = _dev(outputArr, "outputArr", (torch.float64,), 2)
    r = _radem3(radem)
    c = _dev(chiArr, "chiArr", _SFX, 1)
    _same(x, c, "inputArr", "chiArr")
    if o.shape[0] != x.shape[0]:
        raise RuntimeError("no datapoints")
    check(getattr(lib, "xgpr_b200_rbf_feature_gen" + _SFX[x.dtype])(
        x.data_ptr(), x.shape[0], x.shape[1], o.data_ptr(), o.shape[1], r.data_ptr(), r.shape[2],
        c.data_ptr(), c.shape[0], int(bool(fitIntercept)), _stream()))
    return 0


def rbf_zv(inputArr, radem, chiArr, fitIntercept, v, t):
    """t (n) += Z v with Z = cudaRBFFeatureGen(inputArr, ...) (column 0 = 1 when fitIntercept), Z never stored."""
    x = _dev(inputArr, "inputArr", _SFX, 2)
    r = _radem3(radem)
    c = _dev(chiArr, "chiArr", _SFX, 1)
    _same(x, c, "inputArr", "chiArr")
    v = _dev(v, "v", (torch.float64,), 1)
    t = _dev(t, "t", (torch.float64,), 1)
    if v.shape[0] != 2 * c.shape[0] or t.shape[0] != x.shape[0]:
        raise RuntimeError("wrong array sizes")
    check(getattr(lib, "xgpr_b200_rbf_zv" + _SFX[x.dtype])(
        x.data_ptr(), x.shape[0], x.shape[1], r.data_ptr(), r.shape[2], c.data_ptr(), c.shape[0],
        int(bool(fitIntercept)), v.data_ptr(), t.data_ptr(), _stream()))


def rbf_ztv(inputArr, radem, chiArr, fitIntercept, t, w):
    """w (2F) += Z^T t, Z as in rbf_zv."""
    x = _dev(inputArr, "inputArr", _SFX, 2)
    r = _radem3(radem)
    c = _dev(chiArr, "chiArr", _SFX, 1)
    _same(x, c, "inputArr", "chiArr")
    t = _dev(t, "t", (torch.float64,), 1)
    w = _dev(w, "w", (torch.float64,), 1)
    if w.shape[0] != 2 * c.shape[0] or t.shape[0] != x.shape[0]:
        raise RuntimeError("wrong array sizes")
    check(getattr(lib, "xgpr_b200_rbf_ztv" + _SFX[x.dtype])(
        x.data_ptr(), x.shape[0], x.shape[1], r.data_ptr(), r.shape[2], c.data_ptr(), c.shape[0],
        int(bool(fitIntercept)), t.data_ptr(), w.data_ptr(), _stream()))


def rbf_zzv_supported(xdim_last, num_freqs):
    """True when the one-pass fused matvec can hold all repeats of a row in one CTA."""
    return bool(lib.xgpr_b200_rbf_zzv_supported(int(xdim_last), int(num_freqs)))


def rbf_zzv(inputArr, radem, chiArr, fitIntercept, v, w):
    """w (2F) += Z^T (Z v) in ONE pass over the rows (t reduced on chip), Z as in rbf_zv."""
    x = _dev(inputArr, "inputArr", _SFX, 2)
    r = _radem3(radem)
    c = _dev(chiArr, "chiArr", _SFX, 1)
    _same(x, c, "inputArr", "chiArr")
    v = _dev(v, "v", (torch.float64,), 1)
    w = _dev(w, "w", (torch.float64,), 1)
    if w.shape[0] != 2 * c.shape[0] or v.shape[0] != 2 * c.shape[0]:
        raise RuntimeError("wrong array sizes")
    check(getattr(lib, "xgpr_b200_rbf_zzv" + _SFX[x.dtype])(
        x.data_ptr(), x.shape[0], x.shape[1], r.data_ptr(), r.shape[2], c.data_ptr(), c.shape[0],
        int(bool(fitIntercept)), v.data_ptr(), w.data_ptr(), _stream()))


def cudaRBFFeatureGenC32(inputArr, outputArr, radem, chiArr):
    """outputArr (float32, n x 2F) <- the unscaled cos / sin values of cudaRBFFeatureGen: the lossless resident encoding."""
    x = _dev(inputArr, "inputArr", (torch.float32,), 2)
    o = _dev(outputArr, "outputArr", (torch.float32,), 2)
    r = _radem3(radem)
    c = _dev(chiArr, "chiArr", (torch.float32,), 1)
    if o.shape[0] != x.shape[0]:
        raise RuntimeError("no datapoints")
    check(lib.xgpr_b200_rbf_feature_gen_c32(x.data_ptr(), x.shape[0], x.shape[1], o.data_ptr(), o.shape[1],
                                            r.data_ptr(), r.shape[2], c.data_ptr(), c.shape[0], _stream()))
    return 0


def unscale_to_f32(z, scale, out):
    """out (float32, same shape) = float(z / scale): the lossless 4-byte encoding of sine-cosine features."""
    z = _dev(z, "z", (torch.float64,), 2)
    out = _dev(out, "out", (torch.float32,), 2)
    if out.shape != z.shape:
        raise RuntimeError("wrong array sizes")
    check(lib.xgpr_b200_unscale_to_f32(z.data_ptr(), z.numel(), float(scale), out.data_ptr(), _stream()))
    return out


def cres_zzv_supported(num_rffs):
    return bool(lib.xgpr_b200_cres_zzv_supported(int(num_rffs)))


def cres_zzv(c, vs, w, scale, fitIntercept, skip=None):
    """w (m) += Z^T (Z v) in one pass over the resident float32 encoding c of Z (Z = double(c) * scale, column 0 = 1 when
    fitIntercept); vs = v * scale with vs[0] = v[0] when fitIntercept.  skip: optional device double; the launch is a
    no-op when it is non-zero (the convergence flag of the stream-ordered CG, cg_state()[5:6])."""
    c = _dev(c, "c", (torch.float32,), 2)
    vs = _dev(vs, "vs", (torch.float64,), 1)
    w = _dev(w, "w", (torch.float64,), 1)
    if vs.shape[0] != c.shape[1] or w.shape[0] != c.shape[1]:
        raise RuntimeError("wrong array sizes")
    if skip is None:
        check(lib.xgpr_b200_cres_zzv_f32(c.data_ptr(), c.shape[0], c.shape[1], vs.data_ptr(), w.data_ptr(), float(scale),
                                         int(bool(fitIntercept)), _stream()))
        return
    skip = _dev(skip, "skip", (torch.float64,), 1)
    check(lib.xgpr_b200_cres_zzv_gated_f32(c.data_ptr(), c.shape[0], c.shape[1], vs.data_ptr(), w.data_ptr(),
                                           float(scale), int(bool(fitIntercept)), skip.data_ptr(), _stream()))


CG_STATE_LEN = 8          # [r.z, ||r0||, tol, alpha, beta, done, iterations when done, iteration counter]


def cg_update_a(w, p, x, r, lambda2, state, hist):
    """First half of a device-resident CG iteration (include/xgpr_b200.h): alpha, x += alpha p, residual history,
    r -= alpha (w + lambda2 p), convergence flag.  All arguments are fp64 CUDA vectors; nothing synchronises."""
    for nm, t in (("w", w), ("p", p), ("x", x), ("r", r), ("state", state), ("hist", hist)):
        _dev(t, nm, (torch.float64,), 1)
    m = w.shape[0]
    if p.shape[0] != m or x.shape[0] != m or r.shape[0] != m or state.shape[0] < CG_STATE_LEN:
        raise RuntimeError("wrong array sizes")
    check(lib.xgpr_b200_cg_update_a_f64(w.data_ptr(), p.data_ptr(), x.data_ptr(), r.data_ptr(), m, float(lambda2),
                                        state.data_ptr(), hist.data_ptr(), hist.shape[0], _stream()))


def cg_update_b(r, z, p, state, ps=None, scale=1.0, fitIntercept=False):
    """Second half: beta = r.z / (r.z)_old, p = z + beta p, and (optionally) ps = p * scale with ps[0] = p[0] when
    fitIntercept -- the scaled direction cres_zzv consumes."""
    for nm, t in (("r", r), ("z", z), ("p", p), ("state", state)):
        _dev(t, nm, (torch.float64,), 1)
    m = r.shape[0]
    if z.shape[0] != m or p.shape[0] != m or state.shape[0] < CG_STATE_LEN:
        raise RuntimeError("wrong array sizes")
    psp = 0
    if ps is not None:
        _dev(ps, "ps", (torch.float64,), 1)
        if ps.shape[0] != m:
            raise RuntimeError("wrong array sizes")
        psp = ps.data_ptr()
    check(lib.xgpr_b200_cg_update_b_f64(r.data_ptr(), z.data_ptr(), p.data_ptr(), psp, m, float(scale),
                                        int(bool(fitIntercept)), state.data_ptr(), _stream()))


def cudaRBFGrad(inputArr, outputArr, gradArr, radem, chiArr, sigma, fitIntercept):
    x = _dev(inputArr, "inputArr", _SFX, 2)
    o = _dev(outputArr, "outputArr", (torch.float64,), 2)
    g = _dev(gradArr, "gradArr", (torch.float64,), 3)
    r = _radem3(radem)
    c = _dev(chiArr, "chiArr", _SFX, 1)
    _same(x, c, "inputArr", "chiArr")
    if o.shape[0] != x.shape[0]:
        raise RuntimeError("no datapoints")
    if g.shape[0] != o.shape[0] or g.shape[1] != o.shape[1] or g.shape[2] != 1:
        raise RuntimeError("Wrong array sizes.")
    check(getattr(lib, "xgpr_b200_rbf_grad" + _SFX[x.dtype])(
        x.data_ptr(), x.shape[0], x.shape[1], o.data_ptr(), g.data_ptr(), o.shape[1], r.data_ptr(), r.shape[2],
        c.data_ptr(), c.shape[0], float(sigma), int(bool(fitIntercept)), _stream()))
    return 0


def cudaMiniARDGrad(inputArr, outputArr, precompWeights, sigmaMap, sigmaVals, gradArr, fitIntercept, overwrite=False):
    """overwrite=True (an addition to the reference signature): outputArr / gradArr are written without being read, i.e.
    the values a zero-filled gradArr would receive -- they may be torch.empty storage (include/xgpr_b200.h)."""
    x = _dev(inputArr, "inputArr", _SFX, 2)
    o = _dev(outputArr, "outputArr", (torch.float64,), 2)
    w = _dev(precompWeights, "precompWeights", _SFX, 2)
    sm = _dev(sigmaMap, "sigmaMap", (torch.int32,), 1)
    sv = _dev(sigmaVals, "sigmaVals", (torch.float64,), 1)
    g = _dev(gradArr, "gradArr", (torch.float64,), 3)
    _same(x, w, "inputArr", "precompWeights")
    # gpu_rf_gen/rbf_ops/ard_ops.cu:110-126 == cpu_rf_gen/rbf_ops/ard_ops.cpp:66-82
    if x.shape[0] == 0 or o.shape[0] != x.shape[0]:
        raise RuntimeError("no datapoints")
    if g.shape[0] != o.shape[0] or g.shape[1] != o.shape[1]:
        raise RuntimeError("Wrong array sizes.")
    if w.shape[1] != x.shape[1] or o.shape[1] != 2 * w.shape[0] or sm.shape[0] != w.shape[1] \
            or sv.shape[0] != sm.shape[0]:
        raise RuntimeError("Wrong array sizes.")
    check(getattr(lib, ("xgpr_b200_ard_grad_overwrite" if overwrite else "xgpr_b200_ard_grad") + _SFX[x.dtype])(
        x.data_ptr(), x.shape[0], x.shape[1], o.data_ptr(), w.data_ptr(), w.shape[0], sm.data_ptr(),
        sv.data_ptr(), g.data_ptr(), g.shape[2], int(bool(fitIntercept)), _stream()))
    return 0


def _conv_common(inputArr, outputArr, radem, chiArr, seqlengths, odtype):
    x = _dev(inputArr, "inputArr", _SFX, 3)
    o = _dev(outputArr, "outputArr", (odtype,), 2)
    r = _radem3(radem)
    c = _dev(chiArr, "chiArr", _SFX, 1)
    _same(x, c, "inputArr", "chiArr")
    s = _host_i32(seqlengths, "seqlengths")
    if x.shape[0] == 0 or o.shape[0] != x.shape[0]:
        raise RuntimeError("no datapoints")
    if s.shape[0] != x.shape[0]:
        raise RuntimeError("wrong array sizes")
    return x, o, r, c, s


def cudaConv1dMaxpool(inputArr, outputArr, radem, chiArr, seqlengths, convWidth):
    x, o, r, c, s = _conv_common(inputArr, outputArr, radem, chiArr, seqlengths, torch.float32)
    check(getattr(lib, "xgpr_b200_conv1d_maxpool" + _SFX[x.dtype])(
        x.data_ptr(), x.shape[0], x.shape[1], x.shape[2], o.data_ptr(), o.shape[1], r.data_ptr(), r.shape[2],
        c.data_ptr(), c.shape[0], s.ctypes.data, int(convWidth), _stream()))
    return 0


def cudaConv1dFGen(inputArr, outputArr, radem, chiArr, seqlengths, convWidth, scalingType):
    x, o, r, c, s = _conv_common(inputArr, outputArr, radem, chiArr, seqlengths, torch.float64)
    check(getattr(lib, "xgpr_b200_conv1d_fgen" + _SFX[x.dtype])(
        x.data_ptr(), x.shape[0], x.shape[1], x.shape[2], o.data_ptr(), o.shape[1], r.data_ptr(), r.shape[2],
        c.data_ptr(), c.shape[0], s.ctypes.data, int(convWidth), int(scalingType), _stream()))
    return 0


def cudaConvGrad(inputArr, outputArr, radem, chiArr, seqlengths, gradArr, sigma, convWidth, scalingType):
    x, o, r, c, s = _conv_common(inputArr, outputArr, radem, chiArr, seqlengths, torch.float64)
    g = _dev(gradArr, "gradArr", (torch.float64,), 3)
    if g.shape[0] != o.shape[0] or g.shape[1] != o.shape[1] or g.shape[2] != 1:
        raise RuntimeError("wrong array sizes")
    check(getattr(lib, "xgpr_b200_conv1d_grad" + _SFX[x.dtype])(
        x.data_ptr(), x.shape[0], x.shape[1], x.shape[2], o.data_ptr(), o.shape[1], r.data_ptr(), r.shape[2],
        c.data_ptr(), c.shape[0], s.ctypes.data, g.data_ptr(), float(sigma), int(convWidth), int(scalingType),
        _stream()))
    return 0


# ---- downstream consumers (no counterpart in the reference extension: they replace `@` expressions)
def zty_accum(z, y, zty):
    z = _dev(z, "z", (torch.float64,), 2)
    y = _dev(y, "y", (torch.float64,), 1)
    zty = _dev(zty, "zty", (torch.float64,), 1)
    if y.shape[0] != z.shape[0] or zty.shape[0] != z.shape[1]:
        raise RuntimeError("wrong array sizes")
    check(lib.xgpr_b200_zty_accum_f64(z.data_ptr(), z.shape[0], z.shape[1], y.data_ptr(), zty.data_ptr(), _stream()))


def ztz_accum(z, ztz):
    z = _dev(z, "z", (torch.float64,), 2)
    ztz = _dev(ztz, "ztz", (torch.float64,), 2)
    if ztz.shape[0] != z.shape[1] or ztz.shape[1] != z.shape[1]:
        raise RuntimeError("wrong array sizes")
    check(lib.xgpr_b200_ztz_accum_f64(z.data_ptr(), z.shape[0], z.shape[1], ztz.data_ptr(), _stream()))


_TC_WORKSPACE = {}


def _tc_workspace(device, need):
    """Grow-only scratch per (device, stream): no allocation in steady state."""
    key = (device.index, torch.cuda.current_stream().cuda_stream)
    ws = _TC_WORKSPACE.get(key)
    if ws is None or ws.numel() < need:
        _TC_WORKSPACE[key] = None
        ws = torch.empty(need, dtype=torch.uint8, device=device)
        _TC_WORKSPACE[key] = ws
    return ws


def atb_accum_tc(a, z, acc, num_slices=5, digit_bits=8):
    """acc (r, m) += a^T z on the tensor cores (same split-int8 scheme as ztz_accum_tc)."""
    a = _dev(a.contiguous(), "a", (torch.float64,), 2)
    z = _dev(z, "z", (torch.float64,), 2)
    acc = _dev(acc, "acc", (torch.float64,), 2)
    if a.shape[0] != z.shape[0] or acc.shape[0] != a.shape[1] or acc.shape[1] != z.shape[1]:
        raise RuntimeError("wrong array sizes")
    need = lib.xgpr_b200_atb_tc_workspace_bytes(z.shape[0], a.shape[1], z.shape[1], int(num_slices))
    if need < 0:
        raise RuntimeError("num_slices must be in [2, 7]")
    ws = _tc_workspace(z.device, need)
    check(lib.xgpr_b200_atb_accum_tc_f64(a.data_ptr(), z.data_ptr(), z.shape[0], a.shape[1], z.shape[1],
                                         acc.data_ptr(), int(num_slices), int(digit_bits), ws.data_ptr(), ws.numel(),
                                         _stream()))


def ztzq_accum_tc(z, q, acc, num_slices=5, digit_bits=8):
    """acc (m, r) += z^T (z q) on the tensor cores: the second Nystrom pass (rand_nys_constructors.py:18-36)."""
    z = _dev(z, "z", (torch.float64,), 2)
    q = _dev(q.contiguous(), "q", (torch.float64,), 2)
    acc = _dev(acc, "acc", (torch.float64,), 2)
    if q.shape[0] != z.shape[1] or acc.shape[0] != z.shape[1] or acc.shape[1] != q.shape[1]:
        raise RuntimeError("wrong array sizes")
    need = lib.xgpr_b200_ztzq_tc_workspace_bytes(z.shape[0], z.shape[1], q.shape[1], int(num_slices))
    if need < 0:
        raise RuntimeError("num_slices must be in [2, 7] and at most 65,536 features")
    ws = _tc_workspace(z.device, need)
    check(lib.xgpr_b200_ztzq_accum_tc_f64(z.data_ptr(), z.shape[0], z.shape[1], q.data_ptr(), q.shape[1],
                                          acc.data_ptr(), int(num_slices), int(digit_bits), ws.data_ptr(), ws.numel(),
                                          _stream()))


def ztz_accum_tc(z, ztz, num_slices=5, mirror=True, digit_bits=8):
    """Upper triangle of ztz += z^T z on the tensor cores (tcgen05 kind::i8 over `num_slices` exact int8 digit planes of
    `digit_bits` bits, int32 accumulation in TMEM; csrc/gram_tc_kernels.cu).  Default 5 x 8 bits: element error
    <= n * 2^-37 colmax_i colmax_j (typically 4e-13 per unit column scale).  mirror=True also refreshes the lower triangle."""
    z = _dev(z, "z", (torch.float64,), 2)
    ztz = _dev(ztz, "ztz", (torch.float64,), 2)
    if ztz.shape[0] != z.shape[1] or ztz.shape[1] != z.shape[1]:
        raise RuntimeError("wrong array sizes")
    need = lib.xgpr_b200_ztz_tc_workspace_bytes(z.shape[0], z.shape[1], int(num_slices))
    if need < 0:
        raise RuntimeError("num_slices must be in [2, 7]")
    ws = _tc_workspace(z.device, need)
    check(lib.xgpr_b200_ztz_accum_tc_f64(z.data_ptr(), z.shape[0], z.shape[1], ztz.data_ptr(), int(num_slices),
                                         int(digit_bits), 1 if mirror else 0, ws.data_ptr(), ws.numel(), _stream()))


def tc_profile(enable):
    """Bracket every tensor-core GEMM launch with CUDA events (measurement aid, see tc_last_gemm_ms)."""
    lib.xgpr_b200_tc_profile(1 if enable else 0)


def tc_last_gemm_ms():
    return float(lib.xgpr_b200_tc_last_gemm_ms())


def symmetrize_upper(g):
    g = _dev(g, "g", (torch.float64,), 2)
    if g.shape[0] != g.shape[1]:
        raise RuntimeError("wrong array sizes")
    check(lib.xgpr_b200_symmetrize_upper_f64(g.data_ptr(), g.shape[0], _stream()))


def atb_accum(a, z, acc):
    """acc (r, m) += a^T z ; a (n, r) may be a strided view (it is made contiguous), z (n, m)."""
    a = _dev(a.contiguous(), "a", (torch.float64,), 2)
    z = _dev(z, "z", (torch.float64,), 2)
    acc = _dev(acc, "acc", (torch.float64,), 2)
    if a.shape[0] != z.shape[0] or acc.shape[0] != a.shape[1] or acc.shape[1] != z.shape[1]:
        raise RuntimeError("wrong array sizes")
    check(lib.xgpr_b200_atb_accum_f64(a.data_ptr(), z.data_ptr(), z.shape[0], a.shape[1], z.shape[1],
                                      acc.data_ptr(), _stream()))


def cg_matvec_accum(z, v, w, scratch=None):
    z = _dev(z, "z", (torch.float64,), 2)
    v = _dev(v, "v", (torch.float64,), 2)
    w = _dev(w, "w", (torch.float64,), 2)
    if v.shape[0] != z.shape[1] or w.shape != v.shape:
        raise RuntimeError("wrong array sizes")
    if scratch is None:
        scratch = torch.empty((z.shape[0], v.shape[1]), dtype=torch.float64, device=z.device)
    scratch = _dev(scratch, "scratch", (torch.float64,), 2)
    if scratch.shape[0] < z.shape[0] or scratch.shape[1] != v.shape[1]:
        raise RuntimeError("wrong array sizes")
    check(lib.xgpr_b200_cg_matvec_accum_f64(z.data_ptr(), z.shape[0], z.shape[1], v.data_ptr(), v.shape[1],
                                            w.data_ptr(), scratch.data_ptr(), _stream()))


def zty_accum_c32(c, y, zty, scale, fit_intercept):
    """zty (m) += Z^T y with Z given as its float32 encoding c (Z = double(c) * scale; column 0 of c = 1.0f with an intercept)."""
    c = _dev(c, "c", (torch.float32,), 2)
    y = _dev(y, "y", (torch.float64,), 1)
    zty = _dev(zty, "zty", (torch.float64,), 1)
    if y.shape[0] != c.shape[0] or zty.shape[0] != c.shape[1]:
        raise RuntimeError("wrong array sizes")
    check(lib.xgpr_b200_zty_accum_c32(c.data_ptr(), c.shape[0], c.shape[1], y.data_ptr(), zty.data_ptr(), float(scale),
                                      int(bool(fit_intercept)), _stream()))


def srht_sample_c32(c, radem, cols, out, scale, fit_intercept):
    """srht_sample on Z = double(c) * scale given as the float32 encoding c; out (n, rank) float64."""
    c = _dev(c, "c", (torch.float32,), 2)
    r = _dev(radem, "radem", (torch.int8,), 1)
    k = _dev(cols, "cols", (torch.int32,), 1)
    o = _dev(out, "out", (torch.float64,), 2)
    if o.shape[0] != c.shape[0] or o.shape[1] != k.shape[0] or c.shape[1] > r.shape[0]:
        raise RuntimeError("wrong array sizes")
    check(lib.xgpr_b200_srht_sample_c32(c.data_ptr(), c.shape[0], c.shape[1], r.data_ptr(), r.shape[0], k.data_ptr(),
                                        k.shape[0], o.data_ptr(), float(scale), int(bool(fit_intercept)), _stream()))
    return out


def atb_accum_tc_c32(a, c, acc, scale, fit_intercept, num_slices=5, digit_bits=8):
    """acc (r, m) += a^T Z on the tensor cores with Z given as its float32 encoding c."""
    a = _dev(a.contiguous(), "a", (torch.float64,), 2)
    c = _dev(c, "c", (torch.float32,), 2)
    acc = _dev(acc, "acc", (torch.float64,), 2)
    if a.shape[0] != c.shape[0] or acc.shape[0] != a.shape[1] or acc.shape[1] != c.shape[1]:
        raise RuntimeError("wrong array sizes")
    need = lib.xgpr_b200_atb_tc_workspace_bytes(c.shape[0], a.shape[1], c.shape[1], int(num_slices))
    if need < 0:
        raise RuntimeError("num_slices must be in [2, 7]")
    ws = _tc_workspace(c.device, need)
    check(lib.xgpr_b200_atb_accum_tc_c32(a.data_ptr(), c.data_ptr(), c.shape[0], a.shape[1], c.shape[1], acc.data_ptr(),
                                         float(scale), int(bool(fit_intercept)), int(num_slices), int(digit_bits),
                                         ws.data_ptr(), ws.numel(), _stream()))


def cg_matvec_accum_c32(c, vs, w, scale, fit_intercept, scratch=None):
    """w (m, r) += Z^T (Z v) from the float32 encoding c of Z (Z = double(c) * scale; column 0 of c holds 1.0f with an
    intercept); vs = v with rows scaled like Z's columns.  See include/xgpr_b200.h."""
    c = _dev(c, "c", (torch.float32,), 2)
    vs = _dev(vs, "vs", (torch.float64,), 2)
    w = _dev(w, "w", (torch.float64,), 2)
    if vs.shape[0] != c.shape[1] or w.shape != vs.shape:
        raise RuntimeError("wrong array sizes")
    if scratch is None:
        scratch = torch.empty((c.shape[0], vs.shape[1]), dtype=torch.float64, device=c.device)
    scratch = _dev(scratch, "scratch", (torch.float64,), 2)
    if scratch.shape[0] < c.shape[0] or scratch.shape[1] != vs.shape[1]:
        raise RuntimeError("wrong array sizes")
    check(lib.xgpr_b200_cg_matvec_accum_c32(c.data_ptr(), c.shape[0], c.shape[1], vs.data_ptr(), vs.shape[1],
                                            w.data_ptr(), scratch.data_ptr(), float(scale), int(bool(fit_intercept)),
                                            _stream()))


launch_count = _native.launch_count
