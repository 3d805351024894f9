"""Single-RHS consumers on one fp64 feature chunk: Z^T y and Z^T (Z v).  python tools/prof_zty.py [n] [m]
(XGPR_B200_ZTY_V1=1 selects the 8-byte-load kernels)"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from xgpr_b200 import ext

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
m = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
z = torch.randn((n, m), dtype=torch.float64, device="cuda")
y = torch.randn(n, dtype=torch.float64, device="cuda")
v = torch.randn((m, 1), dtype=torch.float64, device="cuda")
zty = torch.zeros(m, dtype=torch.float64, device="cuda")
w = torch.zeros((m, 1), dtype=torch.float64, device="cuda")
s = torch.empty((n, 1), dtype=torch.float64, device="cuda")


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


t1 = timed(lambda: ext.zty_accum(z, y, zty))
t2 = timed(lambda: ext.cg_matvec_accum(z, v, w, s))
zty.zero_(); ext.zty_accum(z, y, zty)
err = float((zty - z.T @ y).abs().max() / (z.T @ y).abs().max())
print(json.dumps({"n": n, "m": m, "v1": bool(os.environ.get("XGPR_B200_ZTY_V1")), "zty_ms": t1,
                  "zty_GB_per_s": n * m * 8 / t1 / 1e6, "zzv_ms": t2, "zzv_GB_per_s": 2 * n * m * 8 / t2 / 1e6,
                  "zty_rel_err": err}))
