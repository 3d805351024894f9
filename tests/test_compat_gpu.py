"""GPU: the drop-in demonstration of SURVEY.md 8(b), boundary B2.  The reference's OWN model class, imported unmodified
(oracle/_ref/xGPR, staged by `make -C oracle ref`), runs with device="cuda": `cupy` resolves to xgpr_b200.compat's
torch-backed stand-in and `xGPR.xgpr_cuda_rfgen_cpp_ext` to the eight callables of libxgpr_b200.so.  fit(mode="exact"),
fit(mode="cg") with the reference's preconditioner, predict with variance and exact_nmll are compared with the fixtures
the same class produced on device="cpu" (tests/golden/make_golden_fits.py, make_golden_nmll.py).

Tolerances: float64 FHT 1e-6 (north_star); float32 FHT 1e-6 on the weights' max-norm (measured ~3e-7),
and the looser bounds written below for the variance (a pinv of a float32-feature Gram)."""
import pytest

from compat_helpers import import_reference, run_regression_case

pytestmark = pytest.mark.gpu

F64 = dict(w_exact=1e-6, pred_exact=1e-6, pred_var=1e-6, ratio=1e-6, w_cg=1e-6, nmll=1e-8)
F32 = dict(w_exact=2e-6, pred_exact=2e-6, pred_var=1e-3, ratio=1e-3, w_cg=2e-6, nmll=1e-6)


@pytest.mark.parametrize("case", [("rbf_f64", "RBF", 256, {}, True, 32, 64, F64),
                                  ("rbf_f32", "RBF", 256, {}, False, 32, 64, F32),
                                  ("conv_f64", "Conv1dRBF", 192, {"conv_width": 3}, True, 24, 32, F64)],
                         ids=lambda c: c[0])
def test_reference_front_end_runs_on_libxgpr_b200(case):
    import torch
    from xgpr_b200 import ext
    xGPR = import_reference()
    name, kname, m, parms, dp, rank, vrffs, tol = case
    l0 = ext.launch_count()
    out = run_regression_case(xGPR, name, kname, m, parms, dp, rank, vrffs, tol)
    assert ext.launch_count() > l0, "no kernel of libxgpr_b200.so was launched: the drop-in did not route to the library"
    torch.cuda.synchronize()
    print(name, out)


def test_reference_kernel_plugin_on_device_arrays():
    """The kernel-plugin calls the L4 loops make (kernel_baseclass.py:269-378) with the reference's own RBF class on
    device="cuda": transform_x returns a stand-in array whose `.T @` products run on this package's kernels."""
    import numpy as np
    xGPR = import_reference()
    import cupy as cp
    from xGPR.kernels import KERNEL_NAME_TO_CLASS
    rng = np.random.default_rng(0)
    x = rng.standard_normal((500, 40)).astype(np.float32)
    y = rng.standard_normal(500)
    kc = KERNEL_NAME_TO_CLASS["RBF"]((500, 40), 512, random_seed=123, device="cpu")
    kg = KERNEL_NAME_TO_CLASS["RBF"]((500, 40), 512, random_seed=123, device="cuda")
    hp = np.log(np.asarray([0.4, 0.25]))
    kc.set_hyperparams(hp, logspace=True)
    kg.set_hyperparams(hp, logspace=True)
    zc = kc.transform_x(x)
    zg, yg = kg.transform_x_y(x, y)
    assert isinstance(zg, cp.ndarray)
    scale = float(np.abs(zc[:, 1:]).max())
    np.testing.assert_allclose(cp.asnumpy(zg), zc, rtol=1e-5, atol=1e-5 * scale)
    np.testing.assert_allclose(cp.asnumpy(zg.T @ yg), zc.T @ y, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(cp.asnumpy(zg.T @ zg), zc.T @ zc, rtol=1e-5, atol=1e-6)     # tcgen05 Gram behind `@`
    zgc, dzg = kg.gradient_x(x)
    zcc, dzc = kc.gradient_x(x)
    np.testing.assert_allclose(cp.asnumpy(dzg), dzc, rtol=1e-5, atol=1e-5 * float(np.abs(dzc).max()))
