"""Phase timing of the Nystrom preconditioner build (1 GPU): python tools/prof_precond.py [rows]"""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from xgpr_b200 import ext, fitting
from xgpr_b200.kernels import RBF, SRHTCompressor

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 125000
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(123)
x = torch.randn((rows, 64), generator=g, device=dev, dtype=torch.float32)
y = torch.sin(x[:, :8].sum(1)).double()
ds = fitting.build_shard_dataset(x, y, None, 32768)
kern = RBF((rows, 64), 16384, random_seed=123, device="cuda")
kern.set_hyperparams(np.asarray([np.log(np.sqrt(0.0767)), np.log(0.358 / 8 * 4.0)]), logspace=True)
m, rank = 16384, 512


def t(fn, reps=2):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter(); out = fn(); torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return out, best * 1e3


res = {"rows": rows}
comp = SRHTCompressor(rank, m, random_seed=123, device="cuda")
xin, yin, ld = next(iter(ds.get_chunked_data()))
(z, yy), res["featurise_chunk_ms"] = t(lambda: kern.transform_x_y(xin, yin, ld))
res["chunk_rows"] = z.shape[0]
cz, res["srht_chunk_ms"] = t(lambda: comp.transform_x(z))
acc = torch.zeros((rank, m), dtype=torch.float64, device=dev)
_, res["atb_chunk_ms"] = t(lambda: ext.atb_accum(cz, z, acc))
zty = torch.zeros(m, dtype=torch.float64, device=dev)
_, res["zty_chunk_ms"] = t(lambda: ext.zty_accum(z, yy, zty))
c_mat, res["srht_acc_ms"] = t(lambda: comp.transform_x(acc))
(_, s1, v1), res["svd_small_ms"] = t(lambda: torch.linalg.svd(c_mat, full_matrices=False))
s1i = 1.0 / torch.sqrt(s1.clamp(min=1e-14))
b, res["matmul_ms"] = t(lambda: acc.T @ v1.T @ (s1i[:, None] * v1))
_, res["svd_tall_ms"] = t(lambda: torch.linalg.svd(b, full_matrices=False))
(q, r), res["qr_tall_ms"] = t(lambda: torch.linalg.qr(b))
_, res["svd_r_ms"] = t(lambda: torch.linalg.svd(r, full_matrices=False))
_, res["eigh_gram_ms"] = t(lambda: torch.linalg.eigh(b.T @ b))
_, res["precond_total_ms"] = t(lambda: fitting.RandNysPreconditioner(kern, ds, rank, False, 123, "srht"), 1)
print(json.dumps(res))
