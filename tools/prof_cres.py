"""Resident one-pass matvec w += Z^T (Z v) over the float32 encoding: python tools/prof_cres.py [n] [m]"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from xgpr_b200 import ext

n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
m = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
c = torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1)
v = torch.randn(m, dtype=torch.float64, device="cuda")
w = torch.zeros(m, dtype=torch.float64, device="cuda")
scale = (1.0 / (m / 2 - 0.5)) ** 0.5
for _ in range(3):
    ext.cres_zzv(c, v, w, scale, True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    ext.cres_zzv(c, v, w, scale, True)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(json.dumps({"n": n, "m": m, "ms": ms, "GB_per_s": n * m * 4 / ms / 1e6, "rows_per_s": n / ms * 1e3}))
