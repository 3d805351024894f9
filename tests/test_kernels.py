"""Kernel plugin layer (xgpr_b200.kernels) against fixtures generated from the reference's Python
kernel classes (tests/golden/make_golden_kernels.py).

CPU part : host parameter draws are BIT-IDENTICAL to the reference's for the same seed.
GPU part : transform_x / gradient_x on device='cuda' match the reference's CPU outputs.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, load_golden


def _cases():
    # make_golden_kernels imports the reference at module import; read its CASES table without importing it
    import ast
    src = open(os.path.join(GOLDEN, "make_golden_kernels.py")).read()
    tree = ast.parse(src)
    for node in tree.body:
        if isinstance(node, ast.Assign) and getattr(node.targets[0], "id", "") == "CASES":
            return ast.literal_eval(node.value)
    raise AssertionError("CASES not found")


CASES = _cases()
IDS = [c[0] for c in CASES]
PARAM_ATTRS = ("radem_diag", "chi_arr", "radem_diag1", "chi_arr1", "radem_diag2", "chi_arr2",
               "full_ard_weights", "ard_position_key")


def _make(case, device):
    from xgpr_b200.kernels import KERNEL_NAME_TO_CLASS
    name, kname, xdim, num_rffs, parms, hp = case
    kern = KERNEL_NAME_TO_CLASS[kname](tuple(xdim), num_rffs, random_seed=123, device=device,
                                       double_precision=False, kernel_spec_parms=parms)
    kern.set_hyperparams(np.log(np.asarray(hp)), logspace=True)
    return kern


@pytest.mark.parametrize("case", CASES, ids=IDS)
def test_host_draws_bit_identical(case):
    g = load_golden(os.path.join(GOLDEN, f"l2_{case[0]}.npz"))
    kern = _make(case, "cpu")
    for attr in PARAM_ATTRS:
        if attr in g:
            mine = getattr(kern, attr)
            assert mine.dtype == g[attr].dtype, attr
            assert np.array_equal(mine, g[attr]), attr
    assert kern.get_num_rffs() == g["z"].shape[1]
    assert np.array_equal(kern.get_hyperparams(logspace=False), g["hyperparams"])


def test_srht_draws_bit_identical():
    from xgpr_b200.kernels import SRHTCompressor
    g = load_golden(os.path.join(GOLDEN, "l2_srht.npz"))
    comp = SRHTCompressor(24, 100, device="cpu", double_precision=True, random_seed=123)
    assert np.array_equal(comp.radem, g["radem"])
    assert np.array_equal(comp.col_sampler, g["col_sampler"])


def test_registry_and_errors():
    from xgpr_b200.kernels import KERNEL_NAME_TO_CLASS, RBF, Conv1dRBF, MiniARD, Matern
    assert sorted(KERNEL_NAME_TO_CLASS) == sorted(
        ["RBF", "Matern", "Cauchy", "Conv1dRBF", "Conv1dMatern", "Conv1dCauchy", "Conv1dTwoLayer",
         "GraphRBF", "GraphMatern", "GraphCauchy", "Linear", "MiniARD"])
    with pytest.raises(RuntimeError):
        RBF((10, 8), 33)                       # odd num_rffs for a sine-cosine kernel
    with pytest.raises(ValueError):
        RBF((10, 8, 3), 32)                    # 3-d input for a fixed-vector kernel
    with pytest.raises(ValueError):
        Conv1dRBF((10, 8, 3), 32)              # conv_width missing
    with pytest.raises(ValueError):
        Matern((10, 8), 32)                    # nu missing
    with pytest.raises(ValueError):
        MiniARD((10, 8), 32, kernel_spec_parms={"split_points": []})
    k = RBF((10, 8), 32)
    with pytest.raises(RuntimeError):
        k.device = "tpu"
    with pytest.raises(RuntimeError):
        k.transform_x(np.zeros((3, 8)))        # device='cpu': no CPU fallback in this package
    assert k.validate_new_datapoints(np.zeros((3, 8))) and not k.validate_new_datapoints(np.zeros((3, 9)))
    k.set_bounds(np.log(np.asarray([[1e-2, 2.0], [1e-3, 10.0]])))
    assert np.allclose(k.get_bounds(logspace=False), [[1e-2, 2.0], [1e-3, 10.0]])
    with pytest.raises(RuntimeError):
        k.set_bounds(np.zeros((3, 2)))


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=IDS)
def test_transform_and_gradient_match_reference(case):
    g = load_golden(os.path.join(GOLDEN, f"l2_{case[0]}.npz"))
    kern = _make(case, "cuda")
    seqlen = g.get("seqlen")
    z = kern.transform_x(g["x"], seqlen).cpu().numpy()
    scale = float(np.abs(g["z"][:, 1:]).max())
    np.testing.assert_allclose(z, g["z"], rtol=1e-5, atol=1e-5 * scale)
    zg, dz = kern.gradient_x(g["x"], seqlen)
    zg, dz = zg.cpu().numpy(), dz.cpu().numpy()
    assert dz.shape == g["dz"].shape
    np.testing.assert_allclose(zg, g["zg"], rtol=1e-5, atol=1e-5 * scale)
    if dz.size:
        np.testing.assert_allclose(dz, g["dz"], rtol=1e-5, atol=1e-5 * float(np.abs(g["dz"]).max()))
    if "precomputed_weights" in g:
        np.testing.assert_array_equal(kern.precomputed_weights.cpu().numpy(), g["precomputed_weights"])


@pytest.mark.gpu
def test_srht_compressor_matches_reference():
    import torch
    from xgpr_b200.kernels import SRHTCompressor
    g = load_golden(os.path.join(GOLDEN, "l2_srht.npz"))
    comp = SRHTCompressor(24, 100, device="cuda", double_precision=True, random_seed=123)
    x = torch.from_numpy(g["x"]).cuda()
    assert np.array_equal(comp.transform_x(x).cpu().numpy(), g["out"])
    assert np.array_equal(comp.transform_x(x, no_compression=True).cpu().numpy(), g["full"])


@pytest.mark.gpu
def test_transform_accepts_resident_and_pinned_inputs():
    import torch
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(0)
    x = rng.standard_normal((50, 64))
    k = RBF((50, 64), 512, device="cuda")
    k.set_hyperparams(np.log(np.asarray([0.3, 0.2])))
    z0 = k.transform_x(x)
    z1 = k.transform_x(torch.from_numpy(x.astype(np.float32)).pin_memory())
    xd = torch.from_numpy(x.astype(np.float32)).cuda()
    z2 = k.transform_x(xd)
    assert torch.equal(z0, z1) and torch.equal(z0, z2)
    assert torch.equal(xd.cpu(), torch.from_numpy(x.astype(np.float32)))   # caller's tensor untouched


def test_maxpool_feature_extractor_draws_bit_identical():
    """FHTMaxpoolConv1dFeatureExtractor (conv_feature_extractor.py:19-118): same host draws as the reference class."""
    from xgpr_b200.kernels import FHTMaxpoolConv1dFeatureExtractor
    g = load_golden(os.path.join(GOLDEN, "l2_maxpool_extractor.npz"))
    fe = FHTMaxpoolConv1dFeatureExtractor(21, 768, random_seed=123, device="cpu", conv_width=9)
    assert np.array_equal(fe.radem_diag, g["radem"]) and np.array_equal(fe.chi_arr, g["chi"])
    assert fe.chi_arr.dtype == np.float32 and fe.radem_diag.dtype == np.int8
    with pytest.raises(ValueError):
        fe.transform_x(g["x"], None)
    with pytest.raises(RuntimeError):
        fe.transform_x(g["x"], g["seqlen"])              # device="cpu": no CPU fallback here


@pytest.mark.gpu
def test_maxpool_feature_extractor_matches_reference():
    """... and the same features, bit for bit (float32 max-pool of bit-identical pre-activations)."""
    from xgpr_b200.kernels import FHTMaxpoolConv1dFeatureExtractor
    g = load_golden(os.path.join(GOLDEN, "l2_maxpool_extractor.npz"))
    fe = FHTMaxpoolConv1dFeatureExtractor(21, 768, random_seed=123, device="cuda", conv_width=9)
    out = fe.transform_x(g["x"], g["seqlen"])
    assert out.dtype.is_floating_point and tuple(out.shape) == (7, 768)
    assert np.array_equal(out.cpu().numpy(), g["out"])
    with pytest.raises(ValueError):
        fe.transform_x(g["x"][:, :, :5], g["seqlen"])
