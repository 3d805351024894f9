#!/usr/bin/env python
"""cfg2-shaped CG fit (RBF, 64 features, num_rffs=16384, Nystrom preconditioner) on 1..8 GPUs, rows sharded.

    python tools/bench_fit.py --rows 262144 --rank 512
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        tools/bench_fit.py --rows 1000000 --rank 512

--rows is the GLOBAL row count (strong scaling: total work fixed).  Each rank generates its own shard on the device
(seed + rank), keeps it resident, and the only collectives are the fp64 all-reduces of fitting.py.  Prints one JSON
line (rank 0): phase times measured with CUDA events, max over ranks.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xgpr_b200 import ext, fitting  # noqa: E402
from xgpr_b200.kernels import RBF  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=262144)
    ap.add_argument("--rffs", type=int, default=16384)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--rank", type=int, default=512)
    ap.add_argument("--chunk", type=int, default=32768)
    ap.add_argument("--tol", type=float, default=1e-6)
    ap.add_argument("--method", default="srht")
    ap.add_argument("--maxiter", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=1, help="untimed full fits first (allocator, cuSOLVER handles, NCCL)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    comm = fitting.RowShardComm()
    lo, hi = (args.rows * rank) // world, (args.rows * (rank + 1)) // world
    g = torch.Generator(device=dev)
    g.manual_seed(123 + rank)
    x = torch.randn((hi - lo, args.dim), generator=g, device=dev, dtype=torch.float32)
    y = torch.sin(x[:, :8].sum(1)) + 0.1 * torch.randn(hi - lo, generator=g, device=dev)
    ds = fitting.build_shard_dataset(x, y.double(), None, args.chunk, comm)
    kern = RBF((args.rows, args.dim), args.rffs, random_seed=123, device="cuda")
    kern.set_hyperparams(np.asarray([np.log(np.sqrt(0.0767)), np.log(0.358 / np.sqrt(args.dim) * 4.0)]), logspace=True)

    def timed(fn):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = fn()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return out, float(t[0]) * 1e-3

    for _ in range(args.warmup):
        pre = fitting.RandNysPreconditioner(kern, ds, args.rank, False, 123, args.method, comm)
        fitting.cg_fit_lib_internal(kern, ds, args.tol, args.maxiter, pre, False, comm)
        del pre
    comm.n_allreduce = 0
    l0 = ext.launch_count()
    pre, t_pre = timed(lambda: fitting.RandNysPreconditioner(kern, ds, args.rank, False, 123, args.method, comm))
    (w, n_iter, losses), t_cg = timed(lambda: fitting.cg_fit_lib_internal(kern, ds, args.tol, args.maxiter, pre, False, comm))
    if rank == 0:
        print(json.dumps({"workload": f"cfg2-shaped CG fit: RBF {args.rows} x {args.dim}, num_rffs={args.rffs}, "
                          f"rank-{args.rank} {args.method} preconditioner, tol {args.tol}",
                          "n_gpus": world, "rows": args.rows, "precond_s": t_pre, "cg_s": t_cg, "fit_s": t_pre + t_cg,
                          "cg_iters": n_iter, "achieved_ratio": pre.achieved_ratio, "final_err": losses[-1],
                          "rows_per_s_fit": args.rows / (t_pre + t_cg), "allreduces": comm.n_allreduce,
                          "gpu_launches": ext.launch_count() - l0, "w_norm": float(torch.linalg.norm(w)),
                          "scaling": "strong", "warmup_fits": args.warmup}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
