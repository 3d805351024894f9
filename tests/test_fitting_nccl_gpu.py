"""GPU, >= 2 devices: the row-sharded fit over NCCL against the single-GPU fit ON THE SAME ROWS (the parity check the
gloo tests cannot give: they run oracle providers on CPU).  Two processes, one per GPU; each holds half of one seeded
dataset; rank 0 also fits the whole dataset alone on its GPU (a world-size-1 communicator) and compares.

Bound: north_star's 1e-6 on fitted weights; what holds is ~1e-10 (the shards change the fp64 summation order only).
Skipped when fewer than two GPUs are visible (the driver's 1-GPU test box)."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(rank, world, port, q):
    import torch.distributed as dist
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import RBF
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    n, d, m = 40000, 64, 4096
    rng = np.random.default_rng(17)
    x = rng.standard_normal((n, d)).astype(np.float32)
    y = np.sin(x[:, :8].sum(1)) + 0.1 * rng.standard_normal(n)
    hp = np.log(np.asarray([np.sqrt(0.0767), 0.358 / np.sqrt(d) * 4.0]))
    res = {}

    def fit(xs, ys, comm):
        ds = fitting.build_shard_dataset(torch.from_numpy(xs).to(dev), ys, None, 8192, comm)
        kern = RBF((n, d), m, random_seed=123, device="cuda")
        kern.set_hyperparams(hp, logspace=True)
        pre = fitting.RandNysPreconditioner(kern, ds, 256, False, 123, "srht", comm)
        op = fitting.ConjugateGrad(comm)
        w, nit, _ = fitting.cg_fit_lib_internal(kern, ds, 1e-8, 300, pre, False, comm, cg_operator=op)
        we, _, _ = fitting.calc_weights_exact(ds, kern, comm)
        return w.cpu().numpy(), nit, pre.achieved_ratio, we.cpu().numpy(), dict(op.last_fit_info), ds

    lo, hi = (n * rank) // world, (n * (rank + 1)) // world
    comm = fitting.RowShardComm()
    w, nit, ratio, we, info, ds = fit(x[lo:hi], y[lo:hi], comm)
    res.update(w=w, nit=nit, ratio=ratio, we=we, info=info, n_allreduce=comm.n_allreduce, mean=ds.get_ymean(),
               std=ds.get_ystd(), n=ds.get_ndatapoints())
    dist.barrier()
    if rank == 0:
        w1, nit1, ratio1, we1, info1, ds1 = fit(x, y, fitting.RowShardComm(single=True))
        res.update(w1=w1, nit1=nit1, ratio1=ratio1, we1=we1, mean1=ds1.get_ymean(), std1=ds1.get_ystd())
        q.put(res)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_nccl_sharded_fit_equals_single_gpu_fit():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_run, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res["n"] == 40000 and res["n_allreduce"] > 0
    np.testing.assert_allclose(res["mean"], res["mean1"], rtol=1e-12)
    np.testing.assert_allclose(res["std"], res["std1"], rtol=1e-12)
    np.testing.assert_allclose(res["ratio"], res["ratio1"], rtol=1e-6)
    assert abs(res["nit"] - res["nit1"]) <= 1
    assert res["info"]["device_cg"]
    scale = float(np.abs(res["w1"]).max())
    assert float(np.abs(res["w"] - res["w1"]).max()) / scale < 1e-6          # north_star bound on fitted weights
    assert float(np.abs(res["we"] - res["we1"]).max()) / float(np.abs(res["we1"]).max()) < 1e-6
