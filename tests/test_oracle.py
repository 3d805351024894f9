"""CPU tests: the NumPy restatement (oracle/sorf_oracle.py) against the golden vectors generated
from the reference's CPU extension, and against that extension itself when oracle/_ref is built."""
import os

import numpy as np
import pytest

from conftest import golden_files, load_golden, ref_ext
from oracle import sorf_oracle as so


def _dt(path):
    return np.float32 if "_f32_" in os.path.basename(path) else np.float64


def _tol(path):
    # float32: glibc cosf/sinf vs NumPy's SIMD cosf/sinf differ by <= ~2 ulp of the argument-side
    # value; pre-activations are bit-exact (asserted separately in test_fht_golden).
    return dict(rtol=2e-6, atol=2e-6) if _dt(path) == np.float32 else dict(rtol=1e-12, atol=1e-13)


@pytest.mark.parametrize("path", golden_files("l1_fht_"), ids=os.path.basename)
def test_fht_golden(path):
    g = load_golden(path)
    y = so.fwht_(g["x"].copy())
    assert np.array_equal(y, g["fht"])            # bit-exact: same butterfly order
    z = so.srht_(g["x"].copy(), g["radem"])
    assert np.array_equal(z, g["srht"])


@pytest.mark.parametrize("path", golden_files("l1_rbf_"), ids=os.path.basename)
def test_rbf_golden(path):
    g = load_golden(path)
    icpt = bool(g["fit_intercept"])
    out = so.rbf_feature_gen(g["x"], g["radem"], g["chi"], icpt)
    np.testing.assert_allclose(out, g["out"], **_tol(path))
    gout, grad = so.rbf_grad(g["x"], g["radem"], g["chi"], float(g["sigma"]), icpt)
    np.testing.assert_allclose(gout, g["gout"], **_tol(path))
    gt = _tol(path)
    gt["atol"] *= max(1.0, float(np.abs(g["grad"]).max()))
    np.testing.assert_allclose(grad, g["grad"], **gt)


@pytest.mark.parametrize("path", golden_files("l1_conv_"), ids=os.path.basename)
def test_conv_golden(path):
    g = load_golden(path)
    w = int(g["conv_width"])
    tol = _tol(path)
    for st in (0, 1, 2):
        out = so.conv_rbf_feature_gen(g["x"], g["radem"], g["chi"], g["seqlen"], w, st)
        t = dict(tol)
        t["atol"] *= max(1.0, float(np.abs(g[f"out{st}"]).max()))
        np.testing.assert_allclose(out, g[f"out{st}"], **t)
        gout, grad = so.conv_rbf_grad(g["x"], g["radem"], g["chi"], g["seqlen"], float(g["sigma"]), w, st)
        np.testing.assert_allclose(gout, g[f"gout{st}"], **t)
        t["atol"] = tol["atol"] * max(1.0, float(np.abs(g[f"grad{st}"]).max()))
        np.testing.assert_allclose(grad, g[f"grad{st}"], **t)
    mp = so.conv_maxpool(g["x"], g["radem_m"], g["chi_m"], g["seqlen"], w)
    assert mp.dtype == np.float32
    assert np.array_equal(mp, g["maxpool"])       # no transcendental: bit-exact


@pytest.mark.parametrize("path", golden_files("l1_ard_"), ids=os.path.basename)
def test_ard_golden(path):
    g = load_golden(path)
    out, grad = so.ard_grad(g["x"], g["w"], g["smap"], g["svals"], bool(g["fit_intercept"]),
                            num_lengthscales=g["grad"].shape[2])
    np.testing.assert_allclose(out, g["out"], rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(grad, g["grad"], rtol=1e-11, atol=1e-11)


def test_against_ref_extension_random_shapes():
    """Fresh seeded inputs through oracle/_ref (the compiled reference) when it is available."""
    ext = ref_ext()
    if ext is None:
        pytest.skip("oracle/_ref not built (make -C oracle ref needs /root/reference)")
    rng = np.random.default_rng(7)
    for dt in (np.float32, np.float64):
        n, d, f = 9, 33, 150
        p = so.padded_width(d)
        x = rng.standard_normal((n, d)).astype(dt)
        radem = rng.choice(np.asarray([-1, 1], np.int8), size=(3, 1, 3 * p))
        chi = np.sqrt(rng.chisquare(p, f)).astype(dt)
        ref_out = np.zeros((n, 2 * f))
        ext.cpuRBFFeatureGen(x, ref_out, radem, chi, True)
        tol = dict(rtol=2e-6, atol=2e-6) if dt == np.float32 else dict(rtol=1e-12, atol=1e-13)
        np.testing.assert_allclose(so.rbf_feature_gen(x, radem, chi, True), ref_out, **tol)


def test_validation_errors():
    """Error behaviour of the boundary (rbf_ops.cpp:49-62, rbf_convolution.cpp:55-82)."""
    x = np.zeros((3, 8), np.float32)
    radem = np.ones((3, 1, 8), np.int8)
    chi = np.ones(4, np.float32)
    with pytest.raises(RuntimeError):
        so.rbf_feature_gen(x, radem, chi, True, out=np.zeros((3, 7)))
    with pytest.raises(RuntimeError):
        so.rbf_feature_gen(x, radem, np.ones(16, np.float32), True)
    xc = np.zeros((2, 6, 4), np.float32)
    rc = np.ones((3, 1, 16), np.int8)
    with pytest.raises(RuntimeError):
        so.conv_rbf_feature_gen(xc, rc, chi, np.asarray([6, 2], np.int32), 3, 0)
    with pytest.raises(RuntimeError):
        so.conv_rbf_feature_gen(xc, rc, chi, np.asarray([6, 7], np.int32), 3, 0)
