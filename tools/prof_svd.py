"""Timing of the serial factorisations of the Nystrom build and of the featurise call overhead (1 GPU)."""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from xgpr_b200 import ext
from xgpr_b200.kernels import RBF


def t(fn, reps=3):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t0 = time.perf_counter(); out = fn(); torch.cuda.synchronize()
        best = min(best, time.perf_counter() - t0)
    return out, best * 1e3


torch.manual_seed(0)
res = {}
b = torch.randn((16384, 512), dtype=torch.float64, device="cuda") * torch.logspace(0, -3, 512, dtype=torch.float64, device="cuda")
c = b.T @ b
for drv in ("gesvd", "gesvdj", "gesvda"):
    try:
        _, res[f"svd_tall_{drv}_ms"] = t(lambda: torch.linalg.svd(b, full_matrices=False, driver=drv))
        _, res[f"svd_512_{drv}_ms"] = t(lambda: torch.linalg.svd(c, full_matrices=False, driver=drv))
    except Exception as e:
        res[f"svd_{drv}_err"] = str(e)[:80]
_, res["eigh_512_ms"] = t(lambda: torch.linalg.eigh(c))
(q, r), res["qr_tall_ms"] = t(lambda: torch.linalg.qr(b))
x = torch.randn((32768, 64), dtype=torch.float32, device="cuda")
kern = RBF((32768, 64), 16384, random_seed=123, device="cuda")
kern.set_hyperparams(np.log(np.asarray([0.277, 0.2])))
_, res["transform_x_32768_ms"] = t(lambda: kern.transform_x(x))
out = torch.empty((32768, 16384), dtype=torch.float64, device="cuda")
_, res["empty_alloc_ms"] = t(lambda: torch.empty((32768, 16384), dtype=torch.float64, device="cuda"))
xs = kern._stage_input(x)
_, res["featuregen_only_ms"] = t(lambda: ext.cudaRBFFeatureGen(xs, out, kern.radem_diag, kern.chi_arr, True))
_, res["col0_write_ms"] = t(lambda: out[:, 0].fill_(1.0))
z = out
a = torch.randn((32768, 512), dtype=torch.float64, device="cuda")
acc = torch.zeros((512, 16384), dtype=torch.float64, device="cuda")
_, res["atb_f64_ms"] = t(lambda: ext.atb_accum(a, z, acc))
_, res["atb_tc6_ms"] = t(lambda: ext.atb_accum_tc(a, z, acc, num_slices=6))
_, res["atb_tc4_ms"] = t(lambda: ext.atb_accum_tc(a, z, acc, num_slices=4))
print(json.dumps(res))
