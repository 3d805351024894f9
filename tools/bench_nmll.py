#!/usr/bin/env python
"""cfg5-shaped approximate NMLL (MiniARD, 128 features, num_rffs=32768, SLQ with 25 probes) on 1..8 GPUs, rows sharded.

    python tools/bench_nmll.py --rows 250000
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 \
        tools/bench_nmll.py --rows 2000000

--rows is the GLOBAL row count (strong scaling).  Each rank generates its shard on the device (seed + rank) and keeps it
resident; the only collectives are the fp64 all-reduces of fitting.py / scoring.py.  Prints one JSON line (rank 0);
time = CUDA events around the whole approximate_nmll call, max over ranks.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xgpr_b200 import ext, fitting, scoring  # noqa: E402
from xgpr_b200.kernels import MiniARD, RBF  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=250000)
    ap.add_argument("--rffs", type=int, default=32768)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--rank", type=int, default=512)
    ap.add_argument("--probes", type=int, default=25)
    ap.add_argument("--chunk", type=int, default=32768)
    ap.add_argument("--tol", type=float, default=1e-6)
    ap.add_argument("--kernel", default="MiniARD", choices=["MiniARD", "RBF"])
    ap.add_argument("--mode", default="srht_2")
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
    sig = 0.358 / np.sqrt(args.dim) * 4.0
    if args.kernel == "MiniARD":
        kern = MiniARD((args.rows, args.dim), args.rffs, random_seed=123, device="cuda",
                       kernel_spec_parms={"split_points": [args.dim // 2]})
        hp = np.log(np.asarray([np.sqrt(0.0767), sig, sig]))
    else:
        kern = RBF((args.rows, args.dim), args.rffs, random_seed=123, device="cuda")
        hp = np.log(np.asarray([np.sqrt(0.0767), sig]))
    settings = {"max_rank": args.rank, "nsamples": args.probes, "nmll_tol": args.tol, "nmll_iter": 500,
                "preconditioner_mode": args.mode}
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = ext.launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    val = scoring.approximate_nmll(hp.copy(), ds, kern, settings, 123, comm)
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"workload": f"cfg5-shaped approximate NMLL: {args.kernel} {args.rows} x {args.dim}, "
                          f"num_rffs={args.rffs}, {args.probes} probes, rank-{args.rank} {args.mode} preconditioner, "
                          f"tol {args.tol}", "n_gpus": world, "rows": args.rows, "nmll": val,
                          "seconds": float(t[0]) * 1e-3, "rows_per_s": args.rows / (float(t[0]) * 1e-3),
                          "allreduces": comm.n_allreduce, "gpu_launches": ext.launch_count() - l0,
                          "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9, "scaling": "strong",
                          "phases_rank0": dict(scoring.LAST_NMLL_INFO)}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
