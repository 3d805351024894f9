"""Host logic of the row-sharded fit (xgpr_b200/fitting.py) under torch.distributed/gloo, world size 1 and 2,
on CPU: sharding, global y statistics, the three all-reduce points and the CG recurrences -- checked against
fixtures generated from the REFERENCE's own toolkits (tests/golden/make_golden_fits.py).  Features and the
consumers of Z come from test-only oracle providers (tests/fit_helpers.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import GOLDEN, load_golden

CASES = {"rbf_f64": ("RBF", 256, {}, True, 32, 64), "conv_f64": ("Conv1dRBF", 192, {"conv_width": 3}, True, 24, 32)}


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(rank, world, port, name, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from fit_helpers import OracleKernel, OracleOps, OracleSRHT
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import KERNEL_NAME_TO_CLASS
    if world > 1:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    g = load_golden(os.path.join(GOLDEN, f"l3_fit_{name}.npz"))
    kname, m, parms, dp, rank_p, vrffs = CASES[name]
    n = g["x"].shape[0]
    lo, hi = (n * rank) // world, (n * (rank + 1)) // world          # contiguous row blocks
    comm = fitting.RowShardComm()
    seq = g.get("seqlen")
    ds = fitting.build_shard_dataset(g["x"][lo:hi], g["y"][lo:hi], None if seq is None else seq[lo:hi],
                                     chunk_size=97, comm=comm)
    kern = KERNEL_NAME_TO_CLASS[kname](g["x"].shape, m, random_seed=123, device="cpu", double_precision=dp,
                                       kernel_spec_parms=parms)
    kern.set_hyperparams(np.log(g["hyperparams"]), logspace=True)
    ok = OracleKernel(kern)
    res = {"mean": ds.get_ymean(), "std": ds.get_ystd(), "n": ds.get_ndatapoints(), "name": name}
    w, _, _ = fitting.calc_weights_exact(ds, ok, comm, OracleOps)
    res["w_exact"] = w.numpy()
    for method in ("srht", "srht_2"):
        pre = fitting.RandNysPreconditioner(ok, ds, rank_p, False, 123, method, comm, OracleOps, OracleSRHT)
        wcg, nit, _ = fitting.cg_fit_lib_internal(ok, ds, 1e-12, 500, pre, False, comm, OracleOps)
        res[f"ratio_{method}"], res[f"w_cg_{method}"], res[f"niter_{method}"] = pre.achieved_ratio, wcg.numpy(), nit
    res["var"] = fitting.calc_variance_exact(ok, ds, vrffs, comm, OracleOps).numpy()
    # the NMLL call sites (xgpr_b200/scoring.py) on the same shards
    from xgpr_b200 import scoring
    gn = load_golden(os.path.join(GOLDEN, f"l3_nmll_{name}.npz"))
    loghp = np.log(gn["hyperparams"])
    res["nmll_exact"] = scoring.exact_nmll(loghp.copy(), ds, ok, comm, OracleOps)
    res["nmll_grad_value"], res["nmll_grad"] = scoring.exact_nmll_gradient(loghp.copy(), ds, ok, 1, comm, OracleOps)
    settings = {"max_rank": int(gn["approx_max_rank"]), "nsamples": int(gn["approx_nsamples"]),
                "nmll_tol": float(gn["approx_tol"]), "nmll_iter": 500, "preconditioner_mode": "srht_2"}
    res["nmll_approx"] = scoring.approximate_nmll(loghp.copy(), ds, ok, settings, 123, comm, OracleOps, OracleSRHT)
    res["n_allreduce"] = comm.n_allreduce
    if rank == 0:
        q.put(res)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _check(res, g, world):
    assert res["n"] == g["x"].shape[0]
    np.testing.assert_allclose(res["mean"], float(g["trainy_mean"]), rtol=1e-12)
    np.testing.assert_allclose(res["std"], float(g["trainy_std"]), rtol=1e-12)
    np.testing.assert_allclose(res["w_exact"], g["w_exact"], rtol=1e-6, atol=1e-9)
    for method in ("srht", "srht_2"):
        np.testing.assert_allclose(res[f"ratio_{method}"], float(g[f"ratio_{method}"]), rtol=1e-6)
        np.testing.assert_allclose(res[f"w_cg_{method}"], g[f"w_cg_{method}"], rtol=1e-6, atol=1e-9)
        assert abs(res[f"niter_{method}"] - int(g[f"niter_{method}"])) <= 1
    np.testing.assert_allclose(res["var"], g["var"], rtol=1e-6, atol=1e-10)
    assert (res["n_allreduce"] > 0) == (world > 1)
    gn = load_golden(os.path.join(GOLDEN, f"l3_nmll_{res['name']}.npz"))
    np.testing.assert_allclose(res["nmll_exact"], float(gn["nmll_exact"]), rtol=1e-8)
    np.testing.assert_allclose(res["nmll_grad_value"], float(gn["nmll_grad_value"]), rtol=1e-8)
    np.testing.assert_allclose(res["nmll_grad"], gn["nmll_grad"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(res["nmll_approx"], float(gn["nmll_approx"]), rtol=1e-6)


@pytest.mark.parametrize("world", [1, 2])
@pytest.mark.parametrize("name", sorted(CASES))
def test_sharded_fit_matches_reference(name, world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_run, args=(r, world, port, name, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    _check(res, load_golden(os.path.join(GOLDEN, f"l3_fit_{name}.npz")), world)
