"""GPU: the nanobind face of the C ABI (xgpr_b200/csrc/nb_module.cpp -> xgpr_b200/lib/xgpr_cuda_rfgen_cpp_ext.*.so): torch
CUDA tensors go in through DLPack, the same kernels run, the same bits come out as through the ctypes face (xgpr_b200.ext);
dtype / device mismatches raise TypeError before the call (nanobind .noconvert()), shape violations RuntimeError."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "xgpr_b200", "lib")


def _module():
    if LIB not in sys.path:
        sys.path.insert(0, LIB)
    try:
        import xgpr_cuda_rfgen_cpp_ext as m
    except ImportError:
        pytest.skip("the nanobind module was not built (needs the nanobind sources, see build_native.build_nanobind)")
    return m


def test_nanobind_module_matches_ctypes_face():
    from xgpr_b200 import ext
    from xgpr_b200.kernels import RBF, Conv1dRBF
    m = _module()
    assert m.abi_version() == 1
    rng = np.random.default_rng(0)
    k = RBF((100, 64), 2048, random_seed=123, device="cuda")
    x = torch.from_numpy((rng.standard_normal((100, 64)) * 0.2).astype(np.float32)).cuda()
    a = torch.empty((100, 2048), dtype=torch.float64, device="cuda")
    b = torch.empty_like(a)
    l0 = m.launch_count()
    assert m.cudaRBFFeatureGen(x, a, k.radem_diag, k.chi_arr, True) == 0
    assert m.launch_count() == l0 + 1
    ext.cudaRBFFeatureGen(x, b, k.radem_diag, k.chi_arr, True)
    assert torch.equal(a, b)
    # keyword names of the reference module, float64 overload, gradient
    kd = RBF((100, 64), 2048, random_seed=123, device="cuda", double_precision=True)
    xd = x.double()
    ga, gb = torch.empty((100, 2048, 1), dtype=torch.float64, device="cuda"), torch.empty((100, 2048, 1), dtype=torch.float64, device="cuda")
    m.cudaRBFGrad(inputArr=xd, outputArr=a, gradArr=ga, radem=kd.radem_diag, chiArr=kd.chi_arr, sigma=0.7, fitIntercept=False)
    ext.cudaRBFGrad(xd, b, gb, kd.radem_diag, kd.chi_arr, 0.7, False)
    assert torch.equal(a, b) and torch.equal(ga, gb)
    # FWHT in place, SRHT
    f1 = torch.from_numpy(rng.standard_normal((33, 256)).astype(np.float32)).cuda()
    f2 = f1.clone()
    m.cudaFastHadamardTransform2D(f1)
    ext.cudaFastHadamardTransform2D(f2)
    assert torch.equal(f1, f2)
    # convolution: seqlengths is a HOST int32 array, as in the reference
    ck = Conv1dRBF((8, 40, 21), 512, random_seed=123, device="cuda", kernel_spec_parms={"conv_width": 9})
    xc = torch.from_numpy((rng.uniform(0, 1, (8, 40, 21)) * 0.3).astype(np.float32)).cuda()
    lens = rng.integers(9, 41, size=8).astype(np.int32)
    ca = torch.zeros((8, 512), dtype=torch.float64, device="cuda")
    cb = torch.zeros_like(ca)
    m.cudaConv1dFGen(xc, ca, ck.radem_diag, ck.chi_arr, lens, 9, 1)
    ext.cudaConv1dFGen(xc, cb, ck.radem_diag, ck.chi_arr, lens, 9, 1)
    assert torch.equal(ca, cb)
    torch.cuda.synchronize()


def test_nanobind_argument_checks():
    from xgpr_b200.kernels import RBF
    m = _module()
    k = RBF((10, 64), 256, random_seed=123, device="cuda")
    x = torch.zeros((10, 64), dtype=torch.float32, device="cuda")
    out = torch.zeros((10, 256), dtype=torch.float64, device="cuda")
    with pytest.raises(TypeError):
        m.cudaRBFFeatureGen(x.cpu(), out, k.radem_diag, k.chi_arr, True)               # host tensor
    with pytest.raises(TypeError):
        m.cudaRBFFeatureGen(x, out.float(), k.radem_diag, k.chi_arr, True)             # wrong output dtype
    with pytest.raises(TypeError):
        m.cudaRBFFeatureGen(x.double(), out, k.radem_diag, k.chi_arr, True)            # x / chi dtype mismatch
    with pytest.raises(RuntimeError):
        m.cudaRBFFeatureGen(x, torch.zeros((10, 254), dtype=torch.float64, device="cuda"), k.radem_diag, k.chi_arr, True)
