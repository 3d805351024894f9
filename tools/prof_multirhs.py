"""Multi-RHS CG matvec on a materialised chunk: python tools/prof_multirhs.py [n] [m] [r]"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from xgpr_b200 import ext

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
m = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
r = int(sys.argv[3]) if len(sys.argv) > 3 else 26
z = torch.randn((n, m), dtype=torch.float64, device="cuda") * m ** -0.5
v = torch.randn((m, r), dtype=torch.float64, device="cuda")
w = torch.zeros((m, r), dtype=torch.float64, device="cuda")
s = torch.empty((n, r), dtype=torch.float64, device="cuda")
for _ in range(2):
    ext.cg_matvec_accum(z, v, w, s)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); ext.cg_matvec_accum(z, v, w, s); b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b)
ref = z.T @ (z @ v)
w.zero_(); ext.cg_matvec_accum(z, v, w, s)
print(json.dumps({"n": n, "m": m, "r": r, "ms": ms, "tflops_useful": 4.0 * n * m * r / ms / 1e9,
                  "tflops_padded32": 4.0 * n * m * 32 * ((r + 31) // 32) / ms / 1e9,
                  "z_GB_per_s": 2 * z.numel() * 8 * ((r + 31) // 32) / ms / 1e6,
                  "max_rel_err": float(((w - ref).abs().max() / ref.abs().max()))}))
