"""GPU: the tensor-core Gram (xgpr_b200_ztz_accum_tc_f64: tcgen05 kind::i8 over exact int8 digit planes, int32 TMEM
accumulators) against float64 NumPy for `xfeatures.T @ xfeatures` (reference: scoring_toolkit/exact_nmll_calcs.py:68-72).

Bars, stated per test: inputs that are exactly representable in the digit planes -> BIT-EXACT; feature-shaped inputs ->
|error| <= n * 2^(-7 s + 2) * colmax_i * colmax_j for s planes (6 planes: ~1e-12 of the column scale, i.e. the level of
DGEMM's own rounding); fitted weights through the tensor-core Gram -> the north_star's 1e-6 against the reference
fixture (tests/test_fitting_gpu.py runs with gram="tc" by default)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ext():
    from xgpr_b200 import ext
    return ext


@pytest.mark.parametrize("shape", [(128, 256), (1, 8), (300, 200), (1000, 700), (4096, 1024), (777, 2049)])
@pytest.mark.parametrize("slices", [3, 6, 7])
def test_dyadic_inputs_bit_exact(shape, slices):
    """Two-digit dyadic inputs: every product and partial sum is exact in int32 and in float64, so the tensor-core Gram
    must equal the float64 Gram bit for bit, including accumulation over calls and the mirrored lower triangle."""
    n, m = shape
    rng = np.random.default_rng(n * 7 + m)
    z = rng.integers(-8000, 8001, size=(n, m)).astype(np.float64) / 16384.0
    ref = z.T @ z
    ext = _ext()
    zt = torch.from_numpy(z).to(DEV)
    g = torch.zeros((m, m), dtype=torch.float64, device=DEV)
    ext.ztz_accum_tc(zt, g, num_slices=slices, digit_bits=7)
    assert np.array_equal(g.cpu().numpy(), ref)
    ext.ztz_accum_tc(zt, g, num_slices=slices, mirror=False, digit_bits=7)      # accumulates; lower triangle untouched
    got = g.cpu().numpy()
    assert np.array_equal(np.triu(got), np.triu(2 * ref))
    assert np.array_equal(np.tril(got, -1), np.tril(ref, -1))
    ext.symmetrize_upper(g)
    assert np.array_equal(g.cpu().numpy(), 2 * ref)


@pytest.mark.parametrize("shape", [(2000, 2048), (5000, 1000), (333, 4100)])
@pytest.mark.parametrize("slices", [4, 5, 6, 7])
def test_feature_inputs_error_bound(shape, slices):
    """Random-feature shaped Z (float32 cosines * scale, intercept column of ones)."""
    n, m = shape
    rng = np.random.default_rng(1)
    z = np.cos(rng.standard_normal((n, m)) * 3).astype(np.float32).astype(np.float64) * np.sqrt(2.0 / m)
    z[:, 0] = 1.0
    ref = z.T @ z
    ext = _ext()
    g = torch.zeros((m, m), dtype=torch.float64, device=DEV)
    ext.ztz_accum_tc(torch.from_numpy(z).to(DEV), g, num_slices=slices, digit_bits=7)
    cmax = np.abs(z).max(axis=0)
    bound = n * 2.0 ** (-7 * slices + 2) * cmax[:, None] * cmax[None, :] + 1e-15 * n * cmax[:, None] * cmax[None, :]
    err = np.abs(g.cpu().numpy() - ref)
    assert (err <= bound).all(), float((err / bound).max())
    assert np.array_equal(g.cpu().numpy(), g.cpu().numpy().T)


def test_row_passes_and_zero_columns():
    """n above the int32 exactness cap (65536 rows per pass) takes several passes; an all-zero column is harmless."""
    n, m = 70000, 300
    rng = np.random.default_rng(2)
    z = rng.integers(-100, 101, size=(n, m)).astype(np.float64) / 128.0
    z[:, 5] = 0.0
    ext = _ext()
    g = torch.zeros((m, m), dtype=torch.float64, device=DEV)
    ext.ztz_accum_tc(torch.from_numpy(z).to(DEV), g, num_slices=4, digit_bits=7)
    assert np.array_equal(g.cpu().numpy(), z.T @ z)
    # 8-bit digits at full magnitude: every leading digit is +-127, so a weight group's accumulator would pass 2^31 after
    # 133k row-pairs; the kernel must drain it in time (5 planes x 70,000 rows = 350k row-pairs in the heaviest group)
    zz = np.where(rng.random((n, m)) < 0.5, -1.0, 1.0) * (127.0 / 256.0 + 127.0 / 65536.0)
    g.zero_()
    ext.ztz_accum_tc(torch.from_numpy(zz).to(DEV), g, num_slices=5, digit_bits=8)
    assert np.array_equal(g.cpu().numpy(), zz.T @ zz)


def test_matches_dfma_gram_on_large_tile_grid():
    """Multi-wave schedule (more tiles than SMs, shared remainder tiles with atomic adds) vs the DFMA Gram kernel."""
    n, m = 1024, 7000          # 28 x 28 tile grid -> 406 upper tiles on 148 SMs
    torch.manual_seed(3)
    z = torch.cos(torch.randn((n, m), device=DEV, dtype=torch.float64)) * (2.0 / m) ** 0.5
    ext = _ext()
    g0 = torch.zeros((m, m), dtype=torch.float64, device=DEV)
    g1 = torch.zeros((m, m), dtype=torch.float64, device=DEV)
    ext.ztz_accum(z, g0)
    ext.ztz_accum_tc(z, g1, num_slices=7, digit_bits=7)
    scale = float(g0.diagonal().max())
    assert float((g0 - g1).abs().max()) <= 1e-13 * scale * 8


def test_argument_errors():
    ext = _ext()
    z = torch.zeros((4, 8), dtype=torch.float64, device=DEV)
    with pytest.raises(RuntimeError):
        ext.ztz_accum_tc(z, torch.zeros((8, 7), dtype=torch.float64, device=DEV))
    with pytest.raises(RuntimeError):
        ext.ztz_accum_tc(z, torch.zeros((8, 8), dtype=torch.float64, device=DEV), num_slices=9)
    with pytest.raises(RuntimeError):
        ext.ztz_accum_tc(z, torch.zeros((8, 8), dtype=torch.float64, device=DEV), digit_bits=6)
    with pytest.raises(TypeError):
        ext.ztz_accum_tc(z.float(), torch.zeros((8, 8), dtype=torch.float64, device=DEV))


@pytest.mark.parametrize("shape", [(300, 64, 200), (4096, 512, 1024), (1000, 300, 2049), (70000, 16, 40)])
def test_atb_dyadic_bit_exact(shape):
    """acc (r, m) += A^T Z on the tensor cores (rand_nys_constructors.py:119), exactly representable inputs."""
    n, r, m = shape
    rng = np.random.default_rng(n + r + m)
    a = rng.integers(-8000, 8001, size=(n, r)).astype(np.float64) / 16384.0
    z = rng.integers(-100, 101, size=(n, m)).astype(np.float64) / 128.0
    ext = _ext()
    acc = torch.zeros((r, m), dtype=torch.float64, device=DEV)
    ext.atb_accum_tc(torch.from_numpy(a).to(DEV), torch.from_numpy(z).to(DEV), acc, num_slices=4, digit_bits=7)
    assert np.array_equal(acc.cpu().numpy(), a.T @ z)
    ext.atb_accum_tc(torch.from_numpy(a).to(DEV), torch.from_numpy(z).to(DEV), acc, num_slices=4, digit_bits=7)
    assert np.array_equal(acc.cpu().numpy(), 2 * (a.T @ z))


def test_atb_general_error_bound():
    n, r, m = 3000, 512, 4096
    rng = np.random.default_rng(5)
    a = rng.standard_normal((n, r))
    z = np.cos(rng.standard_normal((n, m)) * 3) * np.sqrt(2.0 / m)
    ext = _ext()
    acc = torch.zeros((r, m), dtype=torch.float64, device=DEV)
    ext.atb_accum_tc(torch.from_numpy(a).to(DEV), torch.from_numpy(z).to(DEV), acc, num_slices=6, digit_bits=7)
    bound = n * 2.0 ** (-40) * np.abs(a).max(axis=0)[:, None] * np.abs(z).max(axis=0)[None, :]
    assert (np.abs(acc.cpu().numpy() - a.T @ z) <= bound).all()


@pytest.mark.parametrize("shape", [(128, 256), (1000, 700), (20000, 300), (4096, 1024)])
def test_base256_digits_bit_exact_and_bounded(shape):
    """digit_bits = 8 (balanced base-256 digits, 16,384-row passes): dyadic inputs stay bit-exact, columns of ones / just
    under a power of two (leading-digit guard) included; feature-shaped inputs within n * 2^(3 - 8 s) colmax_i colmax_j."""
    n, m = shape
    rng = np.random.default_rng(n + m)
    z = rng.integers(-8000, 8001, size=(n, m)).astype(np.float64) / 16384.0
    z[:, 0] = 1.0
    z[:, 1] = 0.999
    z[:, 2] = -0.5
    ext = _ext()
    g = torch.zeros((m, m), dtype=torch.float64, device=DEV)
    ext.ztz_accum_tc(torch.from_numpy(z).to(DEV), g, num_slices=5, digit_bits=8)
    zq = z.copy()
    zq[:, 1] = np.round(z[:, 1] * 2.0 ** 38) / 2.0 ** 38          # 0.999 is not dyadic: compare against its 5 x 8-bit rounding
    ref = zq.T @ zq
    got = g.cpu().numpy()
    mask = np.ones((m, m), bool)
    mask[1, :] = mask[:, 1] = False
    assert np.array_equal(got[mask], (z.T @ z)[mask])
    np.testing.assert_allclose(got[1, :], ref[1, :], rtol=0, atol=n * 2.0 ** -36)
    f = np.cos(rng.standard_normal((n, m)) * 3).astype(np.float32).astype(np.float64) * np.sqrt(2.0 / m)
    f[:, 0] = 1.0
    for s_, bits in ((5, 8), (6, 8), (4, 8)):
        g.zero_()
        ext.ztz_accum_tc(torch.from_numpy(f).to(DEV), g, num_slices=s_, digit_bits=bits)
        cmax = np.abs(f).max(axis=0)
        bound = n * (2.0 ** (3 - bits * s_) + 1e-15) * cmax[:, None] * cmax[None, :]
        err = np.abs(g.cpu().numpy() - f.T @ f)
        assert (err <= bound).all(), (s_, bits, float((err / bound).max()))
    acc = torch.zeros((64, m), dtype=torch.float64, device=DEV)
    a = rng.integers(-100, 101, size=(n, 64)).astype(np.float64) / 128.0
    ext.atb_accum_tc(torch.from_numpy(a).to(DEV), torch.from_numpy(z).to(DEV), acc, num_slices=5, digit_bits=8)
    got = acc.cpu().numpy()
    keep = np.ones(m, bool)
    keep[1] = False
    assert np.array_equal(got[:, keep], (a.T @ z)[:, keep])


@pytest.mark.parametrize("shape", [(300, 200, 64), (1000, 384, 70), (2049, 1000, 130)])
def test_ztzq_dyadic_exact_products(shape):
    """Second Nystrom pass acc += Z^T (Z Q) (rand_nys_constructors.py:18-36) on the tensor cores.  One-digit dyadic Z and Q:
    T = Z Q is exact, and its entries (|T| < 2^13, multiples of 2^-8) fit the digit planes, so the whole product equals
    float64 NumPy bit for bit; accumulation over two calls included."""
    n, m, r = shape
    rng = np.random.default_rng(n + m + r)
    z = rng.integers(-16, 17, size=(n, m)).astype(np.float64) / 16.0
    q = rng.integers(-16, 17, size=(m, r)).astype(np.float64) / 16.0
    ref = z.T @ (z @ q)
    ext = _ext()
    acc = torch.zeros((m, r), dtype=torch.float64, device=DEV)
    ext.ztzq_accum_tc(torch.from_numpy(z).to(DEV), torch.from_numpy(q).to(DEV), acc)
    assert np.array_equal(acc.cpu().numpy(), ref)
    ext.ztzq_accum_tc(torch.from_numpy(z).to(DEV), torch.from_numpy(q).to(DEV), acc)
    assert np.array_equal(acc.cpu().numpy(), 2 * ref)


def test_ztzq_general_error_and_matches_matvec():
    """Feature-shaped Z (|z| <= scale, rows of very different magnitude) and an orthonormal Q: error against float64 NumPy
    below 1e-10 of the result's scale (5 x 8-bit planes: ~4e-13 per product term), and agreement with cg_matvec_accum."""
    n, m, r = 5000, 2048, 96
    rng = np.random.default_rng(8)
    z = rng.uniform(-1, 1, size=(n, m)) * (2.0 / m) ** 0.5
    z[::7] *= 1e-3                                            # row exponents differ
    z[:, ::5] *= 4.0                                          # column exponents differ
    q, _ = np.linalg.qr(rng.standard_normal((m, r)))
    ref = z.T @ (z @ q)
    ext = _ext()
    zt, qt = torch.from_numpy(z).to(DEV), torch.from_numpy(q).to(DEV)
    acc = torch.zeros((m, r), dtype=torch.float64, device=DEV)
    ext.ztzq_accum_tc(zt, qt, acc)
    scale = float(np.abs(ref).max())
    assert float(np.abs(acc.cpu().numpy() - ref).max()) < 1e-10 * scale
    w = torch.zeros((m, r), dtype=torch.float64, device=DEV)
    ext.cg_matvec_accum(zt, qt, w, torch.empty((n, r), dtype=torch.float64, device=DEV))
    assert float((w - acc).abs().max()) < 1e-10 * scale


def test_ztzq_argument_errors():
    ext = _ext()
    z = torch.zeros((64, 128), dtype=torch.float64, device=DEV)
    with pytest.raises(RuntimeError):
        ext.ztzq_accum_tc(z, torch.zeros((100, 8), dtype=torch.float64, device=DEV),
                          torch.zeros((128, 8), dtype=torch.float64, device=DEV))
    with pytest.raises(RuntimeError):
        ext.ztzq_accum_tc(z, torch.zeros((128, 8), dtype=torch.float64, device=DEV),
                          torch.zeros((128, 9), dtype=torch.float64, device=DEV))


@pytest.mark.parametrize("bad", [float("nan"), float("inf")])
def test_tc_gram_propagates_non_finite_input(bad):
    """A NaN / Inf in Z must poison the Gram (as the reference's fp64 `@` does, so that the Cholesky behind exact_nmll fails
    and the caller gets DEFAULT_SCORE_IF_PROBLEM) instead of being clamped into a plausible-looking finite matrix."""
    import torch
    from xgpr_b200 import ext
    g = torch.Generator(device="cuda")
    g.manual_seed(3)
    z = torch.randn((1000, 512), dtype=torch.float64, device="cuda", generator=g)
    z[123, 45] = bad
    out = torch.zeros((512, 512), dtype=torch.float64, device="cuda")
    ext.ztz_accum_tc(z, out)
    assert not bool(torch.isfinite(out[45, :]).any()) and not bool(torch.isfinite(out[:, 45]).any())
    keep = torch.ones(512, dtype=torch.bool, device="cuda")
    keep[45] = False
    clean = out[keep][:, keep]
    ref = z[:, keep].T @ z[:, keep]
    assert bool(torch.isfinite(clean).all())
    assert float((clean - ref).abs().max()) < 1e-9 * float(ref.abs().max())
    a = torch.randn((1000, 64), dtype=torch.float64, device="cuda", generator=g)
    acc = torch.zeros((64, 512), dtype=torch.float64, device="cuda")
    ext.atb_accum_tc(a, z, acc)
    assert not bool(torch.isfinite(acc[:, 45]).any()) and bool(torch.isfinite(acc[:, keep]).all())


def test_cta_pair_kernel_in_its_own_process():
    """tc_gram_pair_kernel (cta_group::2: two CTAs of a cluster share each column tile, the leader issues M = 256 MMAs for
    both) is opt-in through XGPR_B200_TC_PAIR=1, read once per process: rerun the bit-exact and error-bound cases of this
    file with it set.  The same cases, so the same bounds."""
    import subprocess
    import sys
    if os.environ.get("XGPR_B200_TC_PAIR"):
        pytest.skip("already running with the pair kernel selected")
    env = dict(os.environ, XGPR_B200_TC_PAIR="1")
    here = os.path.abspath(__file__)
    res = subprocess.run([sys.executable, "-m", "pytest", here, "-x", "-q", "-m", "gpu", "-k",
                          "dyadic or error_bound or row_passes or matches_dfma or base256 or ztzq_general"],
                         env=env, capture_output=True, text=True, timeout=600, cwd=os.path.dirname(os.path.dirname(here)))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-2000:]
    assert " passed" in res.stdout
