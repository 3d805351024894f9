"""GPU: the five BASELINE.json configurations at (or near) their real shapes.

cfg1 runs at FULL size against the oracle (features) and float64 NumPy (exact-fit algebra).  The others are too large for a
CPU oracle, so they are checked at full width (num_rffs, sequence length, node count) and a six-figure row count through
size-independent properties of the feature maps:
  * cos^2 + sin^2 = scale^2 for every frequency of every row (RBF family);
  * rows are independent: a row's features do not depend on the batch it is generated in (bit-exact);
  * sum-pooling is additive over windows / nodes: Z(first L positions) - Z(first L-1) = Z(the last window alone), and a
    graph's features are invariant to node order up to fp64 summation order;
  * MiniARD's forward map equals the RBF map on the per-column rescaled input (bit-exact, same SORF parameters).
Tolerances are written where they are asserted."""
import numpy as np
import pytest
import torch

from oracle import sorf_oracle as so

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_cfg1_full_size_features_and_exact_fit():
    """RBF 20,000 x 512 -> 2,048 RFFs: features vs the oracle (rtol 1e-5 of scale), exact-fit weights and predictions vs
    float64 NumPy on the same features (1e-6, north_star), through the tensor-core Gram."""
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(123)
    x = rng.standard_normal((20000, 512)).astype(np.float32)
    y = np.sin(x[:, :8].sum(1)) + 0.1 * rng.standard_normal(20000)
    hp = np.asarray([np.sqrt(0.0767), 0.358 / np.sqrt(512) * 4.0])
    kern = RBF(x.shape, 2048, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(hp), logspace=True)
    z = kern.transform_x(x).cpu().numpy()
    host = RBF(x.shape, 2048, random_seed=123, device="cpu")
    xs = x.copy()
    xs *= hp[1]
    ref = so.rbf_feature_gen(xs, host.radem_diag, host.chi_arr, True)
    ref[:, 0] = 1.0
    np.testing.assert_allclose(z, ref, rtol=1e-5, atol=1e-5 * float(np.abs(ref[:, 1:]).max()))
    ds = fitting.build_shard_dataset(x, y, None, chunk_size=2000)
    w, _, _ = fitting.calc_weights_exact(ds, kern)
    ys = (y - ds.get_ymean()) / ds.get_ystd()
    gram = z.T @ z
    gram[np.diag_indices_from(gram)] += 2 * hp[0] ** 2            # lambda^2 twice, as the reference's exact fit
    w_ref = np.linalg.solve(gram, z.T @ ys)
    np.testing.assert_allclose(w.cpu().numpy(), w_ref, rtol=1e-6, atol=1e-6 * float(np.abs(w_ref).max()))
    pred = fitting.predict(kern, w, x[:500], None, 250, ds.get_ymean(), ds.get_ystd())
    np.testing.assert_allclose(pred, (z[:500] @ w_ref) * ds.get_ystd() + ds.get_ymean(), rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("cfg", [("cfg2", 262144, 64, 16384), ("cfg5", 65536, 128, 32768)])
def test_rbf_feature_properties_at_full_width(cfg):
    from xgpr_b200.kernels import RBF
    _, n, d, m = cfg
    g = torch.Generator(device=DEV)
    g.manual_seed(11)
    x = torch.randn((n, d), generator=g, device=DEV, dtype=torch.float32)
    kern = RBF((n, d), m, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(np.asarray([0.277, 0.358 / np.sqrt(d) * 4.0])), logspace=True)
    scale2 = 1.0 / (m // 2 - 0.5)
    chunk = 32768
    first = None
    for lo in range(0, n, chunk):
        z = kern.transform_x(x[lo:lo + chunk])
        if first is None:
            first = z[:1024].clone()
        pairs = z.view(z.shape[0], m // 2, 2)
        r2 = (pairs ** 2).sum(dim=2)[:, 1:]                      # frequency 0 holds the intercept column
        # float32 cos/sin: |cos^2 + sin^2 - 1| <= ~4 ulp(float32)
        assert float((r2 / scale2 - 1.0).abs().max()) < 1e-6
        assert bool((z[:, 0] == 1.0).all())
    # row independence: the same rows inside a different batch give the same bits
    again = kern.transform_x(x[:1024])
    assert torch.equal(first, again)


def test_cfg3_sum_pool_additivity_full_length():
    """Conv1dRBF, 500 x 21 one-hot, conv_width 9, 8,192 RFFs: Z(L) - Z(L - 1) equals the features of the last window."""
    from xgpr_b200.kernels import Conv1dRBF
    n, length, c, w, m = 256, 500, 21, 9, 8192
    rng = np.random.default_rng(3)
    idx = rng.integers(0, c, size=(n, length))
    x = np.zeros((n, length, c), dtype=np.float32)
    np.put_along_axis(x, idx[:, :, None], 1.0, axis=2)
    kern = Conv1dRBF(x.shape, m, random_seed=123, device="cuda", kernel_spec_parms={"conv_width": w})
    kern.set_hyperparams(np.log(np.asarray([0.277, 0.3])), logspace=True)
    lens = rng.integers(w + 1, length + 1, size=n).astype(np.int32)
    z_full = kern.transform_x(x, lens)
    z_prev = kern.transform_x(x, lens - 1)
    last = np.stack([x[i, lens[i] - w:lens[i]] for i in range(n)])            # (n, w, c): exactly one window
    z_last = kern.transform_x(last, np.full(n, w, dtype=np.int32))
    diff = (z_full - z_prev)[:, 1:]
    # fp64 pooling of float32 cos/sin: the difference of two ~490-term sums cancels to ~1e-13 of the sum
    scale = float(z_full[:, 1:].abs().max())
    assert float((diff - z_last[:, 1:]).abs().max()) < 1e-12 * scale * 500
    assert float(z_last[:, 1:].abs().max()) > 0


def test_cfg4_graph_node_order_invariance():
    """GraphRBF (conv_width 1), <= 64 nodes x 36 features, 8,192 RFFs: permuting a graph's nodes only reorders an fp64 sum."""
    from xgpr_b200.kernels import GraphRBF
    n, nodes, c, m = 4096, 64, 36, 8192
    rng = np.random.default_rng(4)
    x = rng.uniform(0, 1, size=(n, nodes, c)).astype(np.float32)
    lens = rng.integers(4, nodes + 1, size=n).astype(np.int32)
    for i in range(n):
        x[i, lens[i]:] = 0.0
    kern = GraphRBF(x.shape, m, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(np.asarray([0.277, 0.4])), logspace=True)
    z = kern.transform_x(x, lens)
    xp = x.copy()
    for i in range(n):
        xp[i, :lens[i]] = x[i, rng.permutation(lens[i])]
    zp = kern.transform_x(xp, lens)
    scale = float(z.abs().max())
    assert float((z - zp).abs().max()) < 1e-13 * scale * nodes
    # and a graph is the sum of its single-node graphs (first 64 graphs)
    for i in range(0, 64, 8):
        singles = kern.transform_x(x[i, :lens[i], None, :].copy(), np.ones(lens[i], dtype=np.int32))
        np.testing.assert_allclose(singles[:, 1:].sum(0).cpu().numpy(), z[i, 1:].cpu().numpy(), rtol=0,
                                   atol=1e-13 * scale * nodes)


def test_cfg5_miniard_forward_equals_rbf_on_rescaled_input():
    from xgpr_b200.kernels import MiniARD, RBF
    n, d, m = 16384, 128, 32768
    g = torch.Generator(device=DEV)
    g.manual_seed(5)
    x = torch.randn((n, d), generator=g, device=DEV, dtype=torch.float32)
    ard = MiniARD((n, d), m, random_seed=123, device="cuda", kernel_spec_parms={"split_points": [64]})
    ard.set_hyperparams(np.log(np.asarray([0.277, 0.1, 0.2])), logspace=True)
    rbf = RBF((n, d), m, random_seed=123, device="cuda")
    rbf.set_hyperparams(np.log(np.asarray([0.277, 1.0])), logspace=True)
    wts = torch.cat([torch.full((64,), 0.1, dtype=torch.float64), torch.full((64,), 0.2, dtype=torch.float64)]).to(DEV)
    xr = (x.double() * wts[None, :]).float()                      # one rounding, as NumPy's float32 * float64 product
    z_ard = ard.transform_x(x)
    z_rbf = rbf.transform_x(xr)
    assert torch.equal(z_ard, z_rbf)
