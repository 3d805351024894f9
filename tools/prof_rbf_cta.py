"""RBF feature generation on the D > 1024 path (rbf_cta2_kernel): python tools/prof_rbf_cta.py [n] [d] [m]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from xgpr_b200 import ext
from xgpr_b200.kernels import RBF

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
d = int(sys.argv[2]) if len(sys.argv) > 2 else 3001
m = int(sys.argv[3]) if len(sys.argv) > 3 else 8192
kern = RBF((n, d), m, random_seed=123, device="cuda")
x = torch.randn((n, d), device="cuda", dtype=torch.float32) * 0.02
out = torch.empty((n, m), dtype=torch.float64, device="cuda")
for _ in range(3):
    ext.cudaRBFFeatureGen(x, out, kern.radem_diag, kern.chi_arr, True)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    ext.cudaRBFFeatureGen(x, out, kern.radem_diag, kern.chi_arr, True)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 10
gb = n * (4 * d + 8 * m) / 1e9
print(json.dumps({"n": n, "d": d, "m": m, "lib": os.environ.get("XGPR_B200_LIB", "default"), "ms": ms,
                  "GB_per_s": gb / ms * 1e3, "frac_hbm": gb / ms * 1e3 / 6534.8}))
