"""One-pass fused matvec on cfg2 shape (for ncu): python tools/prof_zzv.py [rows]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from xgpr_b200.kernels import RBF

n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
x = torch.randn((n, 64), device="cuda", dtype=torch.float32)
k = RBF((n, 64), 16384, random_seed=123, device="cuda")
k.set_hyperparams(np.log(np.asarray([0.277, 0.18])), logspace=True)
xs = k.stage_scaled(x)
v = torch.randn(16384, device="cuda", dtype=torch.float64)
w = torch.zeros(16384, device="cuda", dtype=torch.float64)
t = torch.zeros(n, device="cuda", dtype=torch.float64)
for _ in range(2):
    k.fused_zzv(xs, v, w)
    k.fused_zv(xs, v, t)
    k.fused_ztv(xs, t, w)
torch.cuda.synchronize()
for name, fn in (("zzv", lambda: k.fused_zzv(xs, v, w)), ("zv", lambda: k.fused_zv(xs, v, t)), ("ztv", lambda: k.fused_ztv(xs, t, w))):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); fn(); b.record(); torch.cuda.synchronize()
    print(name, "ms", a.elapsed_time(b), "Mrows/s", n / a.elapsed_time(b) / 1e3)
