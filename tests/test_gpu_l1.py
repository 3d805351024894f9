"""GPU parity tests: every C-ABI entry point against (a) the golden vectors generated from the
reference's CPU extension and (b) the NumPy oracle on fresh seeded inputs, incl. config-shaped cases.

Tolerances (float32 FHT): rtol 1e-5 with atol 1e-5 x feature scale; pre-activation
paths without transcendentals (FWHT, SRHT, max-pool) must be BIT-EXACT.
"""
import os

import numpy as np
import pytest
import torch

from conftest import golden_files, load_golden, ref_ext
from oracle import sorf_oracle as so

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _ext():
    from xgpr_b200 import ext
    return ext


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _dt(path):
    return np.float32 if "_f32_" in os.path.basename(path) else np.float64


def _tol(dt, scale=1.0):
    if np.dtype(dt) == np.float32:
        return dict(rtol=1e-5, atol=1e-5 * scale)
    return dict(rtol=1e-9, atol=1e-11 * max(scale, 1.0))


@pytest.mark.parametrize("path", golden_files("l1_fht_"), ids=os.path.basename)
def test_fht_srht_golden(path):
    g = load_golden(path)
    ext = _ext()
    x = T(g["x"])
    ext.cudaFastHadamardTransform2D(x)
    assert np.array_equal(x.cpu().numpy(), g["fht"])
    x = T(g["x"])
    ext.cudaSRHT(x, T(g["radem"]))
    assert np.array_equal(x.cpu().numpy(), g["srht"])


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("shape", [(3001, 1024), (250, 4096), (37, 16384), (5, 32768), (3, 65536), (1000, 64),
                                   (129, 2), (77, 128)])
def test_fht_srht_large(dt, shape):
    """Shapes of tests/fht_operations_tests/test_basic_rfgen.py:26-56 plus the SRHT widths of cfg2/cfg5."""
    rng = np.random.default_rng(123)
    x = rng.standard_normal(shape).astype(dt)
    radem = rng.choice(np.asarray([-1, 1], np.int8), size=shape[1])
    ext = _ext()
    a = T(x)
    ext.cudaFastHadamardTransform2D(a)
    assert np.array_equal(a.cpu().numpy(), so.fwht_(x.copy()))
    a = T(x)
    ext.cudaSRHT(a, T(radem))
    assert np.array_equal(a.cpu().numpy(), so.srht_(x.copy(), radem))


def _run_rbf(x, radem, chi, icpt):
    ext = _ext()
    out = torch.full((x.shape[0], 2 * chi.shape[0]), 7.0, dtype=torch.float64, device=DEV)  # overwritten
    ext.cudaRBFFeatureGen(T(x), out, T(radem), T(chi), icpt)
    return out.cpu().numpy()


def _run_rbf_grad(x, radem, chi, sigma, icpt):
    ext = _ext()
    out = torch.zeros((x.shape[0], 2 * chi.shape[0]), dtype=torch.float64, device=DEV)
    grad = torch.zeros((x.shape[0], 2 * chi.shape[0], 1), dtype=torch.float64, device=DEV)
    ext.cudaRBFGrad(T(x), out, grad, T(radem), T(chi), sigma, icpt)
    return out.cpu().numpy(), grad.cpu().numpy()


@pytest.mark.parametrize("path", golden_files("l1_rbf_"), ids=os.path.basename)
def test_rbf_golden(path):
    g = load_golden(path)
    dt = _dt(path)
    icpt = bool(g["fit_intercept"])
    scale = float(np.abs(g["out"]).max())
    out = _run_rbf(g["x"], g["radem"], g["chi"], icpt)
    np.testing.assert_allclose(out, g["out"], **_tol(dt, scale))
    gout, grad = _run_rbf_grad(g["x"], g["radem"], g["chi"], float(g["sigma"]), icpt)
    np.testing.assert_allclose(gout, g["gout"], **_tol(dt, scale))
    np.testing.assert_allclose(grad, g["grad"], **_tol(dt, float(np.abs(g["grad"]).max())))


# (n, d, num_freqs): cfg1 / cfg2 / cfg5-shaped rows plus the reference's own test shapes
# (tests/fht_operations_tests/test_rbf_rfgen.py:26-68: D=2003, 3001 exercise the >1024 path)
RBF_SHAPES = [(64, 512, 1024), (96, 64, 8192), (40, 128, 16384), (10, 2003, 64), (10, 3001, 2100), (33, 7, 100),
              (300, 20, 50), (17, 1024, 1500), (1, 64, 64), (1000, 84, 2048)]


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("shape", RBF_SHAPES)
def test_rbf_vs_oracle(dt, shape):
    n, d, f = shape
    rng = np.random.default_rng(123)
    p = so.padded_width(d)
    reps = max(1, -(-f // p))
    x = rng.standard_normal((n, d)).astype(dt)
    radem = rng.choice(np.asarray([-1, 1], np.int8), size=(3, 1, reps * p))
    chi = np.sqrt(rng.chisquare(p, f)).astype(dt)
    x = (x * dt(0.5 / np.sqrt(d))).astype(dt)           # sigma*x of O(1), as after `input_x *= sigma`
    ref = so.rbf_feature_gen(x, radem, chi, True)
    ext_ref = ref_ext()
    if ext_ref is not None:                              # the compiled reference itself, when it travelled
        ref2 = np.zeros_like(ref)
        ext_ref.cpuRBFFeatureGen(x, ref2, radem, chi, True)
        ref = ref2
    scale = float(np.abs(ref).max())
    out = _run_rbf(x, radem, chi, True)
    np.testing.assert_allclose(out, ref, **_tol(dt, scale))
    gref, dref = so.rbf_grad(x, radem, chi, 1.234, False)
    gout, grad = _run_rbf_grad(x, radem, chi, 1.234, False)
    np.testing.assert_allclose(gout, gref, **_tol(dt, scale))
    np.testing.assert_allclose(grad, dref, **_tol(dt, float(np.abs(dref).max())))


def test_rbf_large_arguments():
    """|chi * x| up to ~1e4: range reduction must stay accurate (sincosf slow path is not hit here,
    but the Cody-Waite path is stressed)."""
    rng = np.random.default_rng(5)
    n, d, f = 50, 64, 512
    x = (rng.standard_normal((n, d)) * 200).astype(np.float32)
    radem = rng.choice(np.asarray([-1, 1], np.int8), size=(3, 1, 512))
    chi = np.sqrt(rng.chisquare(64, f)).astype(np.float32)
    ref = so.rbf_feature_gen(x, radem, chi, False)
    out = _run_rbf(x, radem, chi, False)
    np.testing.assert_allclose(out, ref, rtol=1e-5, atol=1e-5 * float(np.abs(ref).max()))


@pytest.mark.parametrize("path", golden_files("l1_conv_"), ids=os.path.basename)
def test_conv_golden(path):
    g = load_golden(path)
    dt = _dt(path)
    ext = _ext()
    w = int(g["conv_width"])
    x, radem, chi = T(g["x"]), T(g["radem"]), T(g["chi"])
    n, f = g["x"].shape[0], g["chi"].shape[0]
    for st in (0, 1, 2):
        out = torch.zeros((n, 2 * f), dtype=torch.float64, device=DEV)
        ext.cudaConv1dFGen(x, out, radem, chi, g["seqlen"], w, st)
        np.testing.assert_allclose(out.cpu().numpy(), g[f"out{st}"], **_tol(dt, float(np.abs(g[f"out{st}"]).max())))
        out = torch.zeros((n, 2 * f), dtype=torch.float64, device=DEV)
        grad = torch.zeros((n, 2 * f, 1), dtype=torch.float64, device=DEV)
        ext.cudaConvGrad(x, out, radem, chi, g["seqlen"], grad, float(g["sigma"]), w, st)
        np.testing.assert_allclose(out.cpu().numpy(), g[f"gout{st}"], **_tol(dt, float(np.abs(g[f"gout{st}"]).max())))
        np.testing.assert_allclose(grad.cpu().numpy(), g[f"grad{st}"], **_tol(dt, float(np.abs(g[f"grad{st}"]).max())))
    mp = torch.zeros((n, g["chi_m"].shape[0]), dtype=torch.float32, device=DEV)
    ext.cudaConv1dMaxpool(x, mp, T(g["radem_m"]), T(g["chi_m"]), g["seqlen"], w)
    assert np.array_equal(mp.cpu().numpy(), g["maxpool"])


# (n, m, c, w, f): cfg3-shaped (one-hot 21 channels, w=9), cfg4-shaped (graph, w=1, 36 features), wide window (>1024)
CONV_SHAPES = [(6, 60, 21, 9, 4096), (12, 64, 36, 1, 4096), (4, 12, 130, 9, 700), (9, 33, 8, 4, 48), (5, 20, 3, 2, 5)]


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("shape", CONV_SHAPES)
def test_conv_vs_oracle(dt, shape):
    n, m, c, w, f = shape
    rng = np.random.default_rng(321)
    p = so.padded_width(w * c)
    reps = max(1, -(-f // p))
    if c == 21:
        idx = rng.integers(0, c, size=(n, m))
        x = np.zeros((n, m, c), dt)
        np.put_along_axis(x, idx[:, :, None], 1.0, axis=2)
        x *= dt(0.6)
    else:
        x = (rng.uniform(0, 1, (n, m, c)) * 0.4).astype(dt)
    radem = rng.choice(np.asarray([-1, 1], np.int8), size=(3, 1, reps * p))
    chi = np.sqrt(rng.chisquare(p, f)).astype(dt)
    seqlen = rng.integers(w, m + 1, size=n).astype(np.int32)
    seqlen[0] = m
    ext = _ext()
    for st in (0, 1, 2):
        ref = so.conv_rbf_feature_gen(x, radem, chi, seqlen, w, st)
        out = torch.zeros((n, 2 * f), dtype=torch.float64, device=DEV)
        ext.cudaConv1dFGen(T(x), out, T(radem), T(chi), seqlen, w, st)
        np.testing.assert_allclose(out.cpu().numpy(), ref, **_tol(dt, float(np.abs(ref).max())))
    gref, dref = so.conv_rbf_grad(x, radem, chi, seqlen, 0.77, w, 1)
    out = torch.zeros((n, 2 * f), dtype=torch.float64, device=DEV)
    grad = torch.zeros((n, 2 * f, 1), dtype=torch.float64, device=DEV)
    ext.cudaConvGrad(T(x), out, T(radem), T(chi), seqlen, grad, 0.77, w, 1)
    np.testing.assert_allclose(out.cpu().numpy(), gref, **_tol(dt, float(np.abs(gref).max())))
    np.testing.assert_allclose(grad.cpu().numpy(), dref, **_tol(dt, float(np.abs(dref).max())))
    # max-pool: F == num_rffs, R == reps*P
    fm = reps * p
    chim = np.sqrt(rng.chisquare(p, fm)).astype(dt)
    ref = so.conv_maxpool(x, radem, chim, seqlen, w)
    mp = torch.zeros((n, fm), dtype=torch.float32, device=DEV)
    ext.cudaConv1dMaxpool(T(x), mp, T(radem), T(chim), seqlen, w)
    assert np.array_equal(mp.cpu().numpy(), ref)


def test_conv_accumulates_into_output():
    """`+=` semantics of cudaConv1dFGen (rbf_convolution.cu:144-145): a pre-filled output is added to."""
    rng = np.random.default_rng(3)
    n, m, c, w, f = 4, 10, 4, 3, 16
    x = rng.standard_normal((n, m, c)).astype(np.float32)
    radem = rng.choice(np.asarray([-1, 1], np.int8), size=(3, 1, 16))
    chi = np.sqrt(rng.chisquare(16, f)).astype(np.float32)
    seqlen = np.full(n, m, np.int32)
    ext = _ext()
    out = torch.full((n, 2 * f), 2.5, dtype=torch.float64, device=DEV)
    ext.cudaConv1dFGen(T(x), out, T(radem), T(chi), seqlen, w, 0)
    ref = so.conv_rbf_feature_gen(x, radem, chi, seqlen, w, 0)
    np.testing.assert_allclose(out.cpu().numpy() - 2.5, ref, rtol=1e-5, atol=1e-5 * float(np.abs(ref).max()))


def _ard_tol(dt, ref):
    """fp64: the tensor-core (DMMA) kernel sums the same products as the oracle in a different order -> 1e-9.
    float32: the oracle rounds every product x W to float before widening it (ard_ops.cpp:98); the DMMA kernel keeps the
    exact product, so the argument of cos / sin differs by <= 6e-8 relative per term -- the float32 feature contract
    (rtol 1e-5, atol 1e-5 of scale) applies; the SIMT kernel (XGPR_B200_ARD_SIMT=1) reproduces the rounding exactly."""
    if np.dtype(dt) == np.float64:
        return dict(rtol=1e-9, atol=1e-10)
    return dict(rtol=1e-5, atol=1e-5 * float(np.abs(ref).max()))


@pytest.mark.parametrize("path", golden_files("l1_ard_"), ids=os.path.basename)
def test_ard_golden(path, monkeypatch):
    g = load_golden(path)
    ext = _ext()
    dt = _dt(path)
    n, f, L = g["x"].shape[0], g["w"].shape[0], g["grad"].shape[2]
    for simt in (False, True):
        if simt:
            monkeypatch.setenv("XGPR_B200_ARD_SIMT", "1")
        out = torch.zeros((n, 2 * f), dtype=torch.float64, device=DEV)
        grad = torch.zeros((n, 2 * f, L), dtype=torch.float64, device=DEV)
        ext.cudaMiniARDGrad(T(g["x"]), out, T(g["w"]), T(g["smap"]), T(g["svals"]), grad, bool(g["fit_intercept"]))
        to = dict(rtol=1e-9, atol=1e-11) if simt else _ard_tol(dt, g["out"])
        tg = dict(rtol=1e-9, atol=1e-10) if simt else _ard_tol(dt, g["grad"])
        np.testing.assert_allclose(out.cpu().numpy(), g["out"], **to)
        np.testing.assert_allclose(grad.cpu().numpy(), g["grad"], **tg)


@pytest.mark.parametrize("dt", [np.float32, np.float64])
@pytest.mark.parametrize("case", [(300, 128, 4096, [64, 64]), (77, 200, 500, [150, 50]), (130, 1000, 96, [1, 998, 1]),
                                  (65, 37, 33, [37])])
def test_ard_config_shapes(case, dt):
    """cfg5-shaped (128 inputs, 2 lengthscales) and awkward shapes (odd F, groups that are not multiples of 4, an
    interleaved sigma map, sigma varying INSIDE a group = the general path of the tensor-core kernel)."""
    n, d, f, groups = case
    rng = np.random.default_rng(11)
    x = rng.standard_normal((n, d)).astype(dt)
    w = (rng.standard_normal((f, d)) * 0.3 / np.sqrt(d / 128)).astype(dt)
    smap = np.repeat(np.arange(len(groups), dtype=np.int32), groups)
    if d == 200:
        smap = rng.permutation(smap).astype(np.int32)                 # groups interleaved along the columns
    svals = np.asarray([0.7, 1.3, 0.9])[smap]
    if d == 37:
        svals = rng.uniform(0.5, 1.5, size=d)                         # not constant on the group
    ref_out, ref_grad = so.ard_grad(x, w, smap, svals, True, num_lengthscales=len(groups))
    ext = _ext()
    out = torch.zeros((n, 2 * f), dtype=torch.float64, device=DEV)
    grad = torch.zeros((n, 2 * f, len(groups)), dtype=torch.float64, device=DEV)
    ext.cudaMiniARDGrad(T(x), out, T(w), T(smap), T(svals), grad, True)
    np.testing.assert_allclose(out.cpu().numpy(), ref_out, **_ard_tol(dt, ref_out))
    np.testing.assert_allclose(grad.cpu().numpy(), ref_grad, **_ard_tol(dt, ref_grad))
    # overwrite form: the same bits into storage full of garbage
    out2 = torch.full((n, 2 * f), float("nan"), dtype=torch.float64, device=DEV)
    grad2 = torch.full((n, 2 * f, len(groups)), float("nan"), dtype=torch.float64, device=DEV)
    ext.cudaMiniARDGrad(T(x), out2, T(w), T(smap), T(svals), grad2, True, overwrite=True)
    assert torch.equal(out2, out) and torch.equal(grad2, grad)
    # and the reference contract: a non-zero gradArr is accumulated onto before the rescale (ard_ops.cpp:100,117-121)
    g0 = torch.full_like(grad, 0.25)
    out3 = torch.zeros_like(out)
    ext.cudaMiniARDGrad(T(x), out3, T(w), T(smap), T(svals), g0, True)
    L = len(groups)
    pairs = out3.view(n, f, 2)
    expect = grad.view(n, f, 2, L).clone()
    expect[:, :, 0, :] += -0.25 * pairs[:, :, 1:2]
    expect[:, :, 1, :] += 0.25 * pairs[:, :, 0:1]
    np.testing.assert_allclose(g0.view(n, f, 2, L).cpu().numpy(), expect.cpu().numpy(), rtol=1e-9, atol=1e-12)


def test_error_behaviour():
    """RuntimeError for the shape violations the reference reports (rbf_ops.cu:284-297), TypeError for
    dtype / device mismatches (nanobind noconvert)."""
    ext = _ext()
    x = torch.zeros((3, 8), dtype=torch.float32, device=DEV)
    radem = torch.ones((3, 1, 8), dtype=torch.int8, device=DEV)
    chi = torch.ones(4, dtype=torch.float32, device=DEV)
    with pytest.raises(RuntimeError):
        ext.cudaRBFFeatureGen(x, torch.zeros((3, 7), dtype=torch.float64, device=DEV), radem, chi, True)
    with pytest.raises(RuntimeError):
        ext.cudaRBFFeatureGen(x, torch.zeros((3, 32), dtype=torch.float64, device=DEV), radem,
                              torch.ones(16, dtype=torch.float32, device=DEV), True)
    with pytest.raises(TypeError):
        ext.cudaRBFFeatureGen(x.double(), torch.zeros((3, 8), dtype=torch.float64, device=DEV), radem, chi, True)
    with pytest.raises(TypeError):
        ext.cudaRBFFeatureGen(x, torch.zeros((3, 8), dtype=torch.float32, device=DEV), radem, chi, True)
    xc = torch.zeros((2, 6, 4), dtype=torch.float32, device=DEV)
    rc = torch.ones((3, 1, 16), dtype=torch.int8, device=DEV)
    out = torch.zeros((2, 8), dtype=torch.float64, device=DEV)
    # tests/fht_operations_tests/test_variable_length_seq_handling.py:39-97
    with pytest.raises(RuntimeError):
        ext.cudaConv1dFGen(xc, out, rc, chi, np.asarray([6, 2], np.int32), 3, 0)
    with pytest.raises(RuntimeError):
        ext.cudaConv1dFGen(xc, out, rc, chi, np.asarray([6, 7], np.int32), 3, 0)
    with pytest.raises(TypeError):
        ext.cudaConv1dFGen(xc, out, rc, chi, torch.tensor([6, 6], dtype=torch.int32, device=DEV), 3, 0)
    with pytest.raises(RuntimeError):
        ext.cudaFastHadamardTransform2D(torch.zeros((4, 12), dtype=torch.float32, device=DEV))


@pytest.mark.parametrize("shape", [(500, 256, 1), (2000, 2048, 1), (333, 1000, 26), (4096, 512, 3)])
def test_downstream_consumers(shape):
    """Z^T y, Z^T Z, Z^T (Z v) against float64 NumPy (exact_nmll_calcs.py:68-72, cg_tools.py:42-44)."""
    n, m, r = shape
    rng = np.random.default_rng(0)
    z = rng.standard_normal((n, m)) / np.sqrt(m)
    y = rng.standard_normal(n)
    v = rng.standard_normal((m, r))
    ext = _ext()
    zt = T(z)
    zty = torch.zeros(m, dtype=torch.float64, device=DEV)
    ext.zty_accum(zt, T(y), zty)
    ext.zty_accum(zt, T(y), zty)                       # accumulates
    np.testing.assert_allclose(zty.cpu().numpy(), 2 * (z.T @ y), rtol=1e-11, atol=1e-12)
    ztz = torch.zeros((m, m), dtype=torch.float64, device=DEV)
    ext.ztz_accum(zt, ztz)
    np.testing.assert_allclose(ztz.cpu().numpy(), z.T @ z, rtol=1e-11, atol=1e-12)
    w = torch.zeros((m, r), dtype=torch.float64, device=DEV)
    ext.cg_matvec_accum(zt, T(v), w)
    np.testing.assert_allclose(w.cpu().numpy(), z.T @ (z @ v), rtol=1e-10, atol=1e-11)


@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("shape", [(70, 1500, 100), (33, 8192, 512), (40, 16384, 512), (25, 20000, 700), (9, 1024, 2),
                                   (300, 32768, 1024)])
def test_fused_srht_sample_is_bit_exact(dt, shape):
    """SRHTCompressor.transform_x in one pass (srht_sample_kernel: read the chunk once, keep `rank` columns; a 2-CTA
    cluster with a DSMEM gather at P_M = 32,768) against pad + cudaSRHT + column gather (srht_compressor.py:70-97) --
    the same butterflies, so the same bits -- and against the oracle's SRHT."""
    from xgpr_b200.kernels import SRHTCompressor
    n, m, rank = shape
    ext = _ext()
    rng = np.random.default_rng(2)
    z = rng.standard_normal((n, m)).astype(dt)
    comp = SRHTCompressor(rank, m, device="cuda", double_precision=(dt == np.float64), random_seed=7)
    p = comp.padded_dims
    if not ext.srht_sample_supported(p, torch.float64 if dt == np.float64 else torch.float32):
        pytest.skip("width handled by the three-pass path")
    zt = T(z)
    keep = zt.clone()
    fused = comp.transform_x(zt)
    assert torch.equal(zt, keep), "the fused kernel must not modify its input"
    padded = torch.zeros((n, p), dtype=zt.dtype, device=DEV)
    padded[:, :m] = zt
    ext.cudaSRHT(padded, comp.radem)
    ref = padded[:, comp.truncated_sampler]
    assert fused.shape == (n, rank) and torch.equal(fused, ref)
    zp = np.zeros((n, p), dt)
    zp[:, :m] = z
    assert np.array_equal(fused.cpu().numpy(), so.srht_(zp, comp.radem.cpu().numpy())[:, comp.truncated_sampler.cpu().numpy()])
