"""CPU: host logic of the drop-in layer (xgpr_b200/compat).  The reference front end is imported unmodified with
device="cuda"; the cupy stand-in is pointed at torch CPU tensors and the eight extension callables are served by the
reference's CPU extension (oracle/_ref) through a TEST-ONLY bridge, so that every array expression of the reference's
device="cuda" branches (slice assignment, .flat[::n+1] +=, NumPy-scalar promotion, linalg, solve_triangular ...) is
exercised without a GPU.  The GPU twin (tests/test_compat_gpu.py) runs the same cases on libxgpr_b200.so."""
import numpy as np
import pytest
import torch

from compat_helpers import import_reference, run_regression_case

F64 = dict(w_exact=1e-6, pred_exact=1e-6, pred_var=1e-6, ratio=1e-6, w_cg=1e-6, nmll=1e-8)


@pytest.fixture()
def cpu_bridge(monkeypatch):
    xGPR = import_reference()
    import xgpr_cpu_rfgen_cpp_ext as cpu_ext
    import xgpr_b200.compat.cupy_shim as shim
    from xgpr_b200 import ext
    monkeypatch.setattr(shim, "_device", lambda: torch.device("cpu"))
    monkeypatch.setattr(shim, "_routed_matmul", lambda a, o: None)
    # `x *= np.float64(s)` on float32 goes to the scale_input kernel on the GPU; same arithmetic here
    monkeypatch.setattr(ext, "scale_input_", lambda t, s=1.0, col_weights=None: t.copy_((t.double() * s).to(t.dtype)))

    def np_view(a):
        t = a._t if isinstance(a, shim.ndarray) else a
        return t.numpy() if isinstance(t, torch.Tensor) else t

    def bridge(cpu_name):
        fn = getattr(cpu_ext, cpu_name)
        return lambda *args: fn(*[np_view(a) for a in args])

    # the kernel modules bound the callables at import time (`from xGPR.xgpr_cuda_rfgen_cpp_ext import ...`)
    import xGPR.kernels.basic_kernels.sorf_kernel_baseclass as m1
    import xGPR.kernels.convolution_kernels.conv_kernel_baseclass as m2
    import xGPR.kernels.srht_compressor as m3
    monkeypatch.setattr(m1, "cudaRBFFeatureGen", bridge("cpuRBFFeatureGen"))
    monkeypatch.setattr(m1, "cudaRBFGrad", bridge("cpuRBFGrad"))
    monkeypatch.setattr(m2, "cudaConv1dFGen", bridge("cpuConv1dFGen"))
    monkeypatch.setattr(m2, "cudaConvGrad", bridge("cpuConvGrad"))
    monkeypatch.setattr(m3, "cudaSRHT", bridge("cpuSRHT"))
    return xGPR


@pytest.mark.parametrize("case", [("rbf_f64", "RBF", 256, {}, True, 32, 64), ("conv_f64", "Conv1dRBF", 192,
                                                                              {"conv_width": 3}, True, 24, 32)],
                         ids=lambda c: c[0])
def test_reference_cuda_branches_on_the_stand_in(cpu_bridge, case):
    name, kname, m, parms, dp, rank, vrffs = case
    run_regression_case(cpu_bridge, name, kname, m, parms, dp, rank, vrffs, F64)


def test_stand_in_array_semantics():
    import xgpr_b200.compat.cupy_shim as cp
    saved = cp._device
    cp._device = lambda: torch.device("cpu")
    try:
        a = cp.zeros((4, 4))
        a.flat[::5] += 2.5
        a.flat[::5] -= 0.5
        assert np.allclose(cp.asnumpy(a), 2.0 * np.eye(4))
        x = cp.asarray(np.arange(6, dtype=np.float32).reshape(2, 3))
        assert (x * np.float64(0.1)).dtype == np.float64 and (x * 0.1).dtype == np.float32      # NEP 50
        r = cp.zeros((5, 2, 3))
        r[:, 0, :] = cp.asarray(np.ones((5, 3)))
        alpha = (r[:, 0, :] * r[:, 0, :]).sum(axis=0) / (r[:, 0, :] * 2).sum(axis=0)
        assert alpha.shape == (3,) and np.allclose(cp.asnumpy(alpha), 0.5)
        e = cp.linalg.norm(r[:, 0, :], axis=0)
        assert float(e[0]) == np.sqrt(5.0) and bool(e.max() < 3)
        s = (cp.asarray(np.array([4.0, -1.0])) ** 2 - 2.0).clip(min=0)
        assert np.allclose(cp.asnumpy(s), [14, 0])
        z = cp.asarray(np.array([3.0, 4.0]))
        z[cp.asarray(np.array([1e-20, 1.0])) < 1e-14] = 0
        assert np.allclose(cp.asnumpy(z), [0, 4])
        assert np.allclose(cp.asnumpy(1 - z), [1, -3]) and np.allclose(cp.asnumpy(8 / cp.asarray(np.array([2.0, 4.0]))), [4, 2])
    finally:
        cp._device = saved
