#!/usr/bin/env python
"""Exact-fit pass (features -> Z^T y, Z^T Z -> Cholesky solve) on config-shaped data, one GPU, device-resident inputs.

    python tools/bench_exact.py            # cfg1 (full size), cfg3 and cfg4 (row subsets, full width)

Reports per config: seconds for the accumulation pass with the tensor-core Gram and with the fp64-pipe Gram, the solve, and
rows/s of the whole exact fit.  CUDA events; one warm-up pass."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xgpr_b200 import fitting  # noqa: E402
from xgpr_b200.kernels import RBF, Conv1dRBF, GraphRBF  # noqa: E402

DEV = torch.device("cuda", 0)


def timed(fn):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    out = fn()
    b.record()
    torch.cuda.synchronize()
    return out, a.elapsed_time(b) * 1e-3


def run(name, kern, x, y, seqlen, chunk):
    ds = fitting.build_shard_dataset(x, y, seqlen, chunk)
    res = {"case": name, "rows": int(x.shape[0]), "num_rffs": kern.get_num_rffs(), "chunk": chunk}
    for gram in ("tc", "f64"):
        fitting.calc_design_mat(ds, kern, gram=gram)                       # warm-up
        (ztz, zty, _), t_acc = timed(lambda: fitting.calc_design_mat(ds, kern, gram=gram))
        res[f"accumulate_{gram}_s"] = t_acc
    _, t_feat = timed(lambda: [kern.transform_x(xc, sl) for xc, sl in ds.get_chunked_x_data()][-1])
    res["features_only_s"] = t_feat
    fitting.calc_weights_exact(ds, kern)                                  # warm-up (cuSOLVER handle, workspaces)
    (w, _, _), t_fit = timed(lambda: fitting.calc_weights_exact(ds, kern))
    res["exact_fit_s"] = t_fit
    res["rows_per_s_exact_fit"] = x.shape[0] / t_fit
    print(json.dumps(res), flush=True)


def main():
    g = torch.Generator(device=DEV)
    g.manual_seed(123)
    hp = np.log(np.asarray([0.277, 0.1]))
    # cfg1: RBF 20,000 x 512 -> 2,048
    x = torch.randn((20000, 512), generator=g, device=DEV, dtype=torch.float32)
    y = (torch.sin(x[:, :8].sum(1)) + 0.1 * torch.randn(20000, generator=g, device=DEV)).double()
    k = RBF(x.shape, 2048, random_seed=123, device="cuda")
    k.set_hyperparams(hp, logspace=True)
    run("cfg1 RBF 20000x512 -> 2048 (full size)", k, x, y, None, 20000)
    # cfg3: Conv1dRBF one-hot 500 x 21, w 9 -> 8,192 (16,384 of the 200k sequences)
    n = 16384
    idx = torch.randint(0, 21, (n, 500), device=DEV, generator=g)
    x = torch.nn.functional.one_hot(idx, 21).to(torch.float32).contiguous()
    y = torch.randn(n, generator=g, device=DEV).double()
    k = Conv1dRBF(x.shape, 8192, random_seed=123, device="cuda", kernel_spec_parms={"conv_width": 9})
    k.set_hyperparams(np.log(np.asarray([0.277, 0.3])), logspace=True)
    run("cfg3 Conv1dRBF 16384x500x21 w9 -> 8192 (row subset)", k, x, y, np.full(n, 500, dtype=np.int32), 16384)
    # cfg4: GraphRBF 64 x 36 -> 8,192 (65,536 of the 500k graphs)
    n = 65536
    x = torch.rand((n, 64, 36), generator=g, device=DEV, dtype=torch.float32)
    y = torch.randn(n, generator=g, device=DEV).double()
    k = GraphRBF(x.shape, 8192, random_seed=123, device="cuda")
    k.set_hyperparams(np.log(np.asarray([0.277, 0.4])), logspace=True)
    run("cfg4 GraphRBF 65536x64x36 -> 8192 (row subset)", k, x, y, np.full(n, 64, dtype=np.int32), 32768)


if __name__ == "__main__":
    main()
