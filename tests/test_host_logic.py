"""CPU: host-side logic added in round 2 that needs no GPU -- coalesced work chunks, the kernel rebuild behind the
rank-ratio pass, the host parameter recipe of bench.py's reference arm, the cupy stand-in's promotion rules."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_coalesced_chunks_cover_the_same_rows_in_order():
    from xgpr_b200.fitting import ShardDataset
    rng = np.random.default_rng(0)
    x, y = rng.standard_normal((10007, 5)), rng.standard_normal(10007)
    lens = rng.integers(1, 9, size=10007).astype(np.int32)
    ds = ShardDataset(x, y, lens, chunk_size=2000, trainy_mean=0.5, trainy_std=2.0)
    small = list(ds.get_chunked_data())
    for min_rows in (1, 2000, 4100, 10 ** 6):
        big = list(ds.get_coalesced_data(min_rows))
        assert all(b[0].shape[0] % 2000 == 0 for b in big[:-1])              # whole dataset chunks
        np.testing.assert_array_equal(np.concatenate([b[0] for b in big]), np.concatenate([s[0] for s in small]))
        np.testing.assert_array_equal(np.concatenate([b[1] for b in big]), np.concatenate([s[1] for s in small]))
        np.testing.assert_array_equal(np.concatenate([b[2] for b in big]), lens)
        if min_rows >= 4000:
            assert len(big) < len(small)
    xs = [c[0] for c in ds.get_coalesced_data(4000, with_y=False)]
    assert sum(c.shape[0] for c in xs) == 10007 and len(xs) == 3


def test_with_num_rffs_rebuilds_the_kernel_like_the_reference_setter():
    """model.num_rffs = min(num_rffs, 8192) re-initialises the kernel with the same class, seed, settings, hyperparameters
    and bounds (model_baseclass.py:528-540, :465-474); with_num_rffs must give the draws a fresh constructor gives."""
    from xgpr_b200.kernels import Conv1dRBF, Matern, MiniARD
    k = Matern((10, 20), 512, random_seed=7, device="cpu", kernel_spec_parms={"matern_nu": 1.5})
    k.set_hyperparams(np.log([0.3, 0.2]))
    k.set_bounds(np.log(np.asarray([[1e-2, 2.0], [1e-3, 5.0]])))
    k2 = k.with_num_rffs(128)
    ref = Matern((10, 20), 128, random_seed=7, device="cpu", kernel_spec_parms={"matern_nu": 1.5})
    assert np.array_equal(k2.chi_arr, ref.chi_arr) and np.array_equal(k2.radem_diag, ref.radem_diag)
    np.testing.assert_allclose(k2.hyperparams, [0.3, 0.2])
    np.testing.assert_allclose(k2.bounds, k.bounds)
    assert k.get_num_rffs() == 512 and k2.get_num_rffs() == 128
    c2 = Conv1dRBF((4, 30, 21), 512, device="cpu", kernel_spec_parms={"conv_width": 9, "averaging": "sqrt"}).with_num_rffs(64)
    assert c2.num_rffs == 64 and c2.conv_width == 9 and c2.scaling_type == 1
    a = MiniARD((4, 128), 512, device="cpu", kernel_spec_parms={"split_points": [64]})
    a.set_hyperparams(np.log([0.3, 0.2, 0.1]))
    a2 = a.with_num_rffs(64)
    np.testing.assert_allclose(a2.hyperparams, [0.3, 0.2, 0.1])
    np.testing.assert_allclose(a2.full_ard_weights[:64], 0.2)


def test_bench_reference_arm_draws_match_the_kernel_classes():
    """bench.py's reference arm restates the host draws with NumPy / SciPy only (it must not import the package under
    test); they must be the kernel classes' draws, bit for bit."""
    sys.path.insert(0, ROOT)
    import bench
    from xgpr_b200.kernels import RBF, Conv1dRBF
    k, h = RBF((1, 64), 16384, random_seed=123, device="cpu"), bench.HostParams(64, 16384)
    assert np.array_equal(k.radem_diag, h.radem_diag) and np.array_equal(k.chi_arr, h.chi_arr)
    k = Conv1dRBF((1, 500, 21), 16384, random_seed=123, device="cpu", kernel_spec_parms={"conv_width": 9})
    h = bench.HostParams(189, 16384)
    assert np.array_equal(k.radem_diag, h.radem_diag) and np.array_equal(k.chi_arr, h.chi_arr)
    x, y = bench.fit_dataset(1000, 1, 4)
    xf, yf = bench.fit_dataset(1000, 0, 1)
    np.testing.assert_array_equal(x, xf[250:500])                    # every rank slices ONE global dataset
    np.testing.assert_array_equal(y, yf[250:500])


def test_row_shard_comm_single_and_counters():
    from xgpr_b200.fitting import RowShardComm
    c = RowShardComm(single=True)
    t = torch.ones(3, dtype=torch.float64)
    c.all_reduce_(t)
    assert c.world == 1 and c.n_allreduce == 0 and c.bytes_allreduce == 0 and float(t.sum()) == 3.0


def test_bench_nmll_dataset_is_the_same_at_every_world_size():
    """bench.py's cfg5 dataset is defined in 62,500-row blocks drawn from per-block seeds, so the rows a rank holds do not
    depend on how many ranks there are: the rank slices for N = 2 / 3 / 8 concatenate to the N = 1 dataset (CPU generator
    here; the bench uses the device generator with the same construction)."""
    import importlib.util
    import os
    import torch
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    n = 3 * bench.NMLL_BLOCK + 1234
    dev = torch.device("cpu")
    x1, y1 = bench.nmll_dataset(n, 0, 1, dev)
    assert x1.shape == (n, bench.NMLL_DIM) and y1.shape == (n,)
    for world in (2, 3, 8):
        xs, ys = zip(*(bench.nmll_dataset(n, r, world, dev) for r in range(world)))
        assert torch.equal(torch.cat(xs), x1) and torch.equal(torch.cat(ys), y1)
