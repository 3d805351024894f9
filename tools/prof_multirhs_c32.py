"""Multi-RHS matvec w += Z^T (Z V), r = 26 (SLQ): fp64 Z vs its float32 encoding, 32-column vs 24 + 2 tile.
    python tools/prof_multirhs_c32.py [n] [m] [r]      (XGPR_B200_MATVEC_NOTAIL=1 for the 32-column tile)"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from xgpr_b200 import ext

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
m = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
r = int(sys.argv[3]) if len(sys.argv) > 3 else 26
c = torch.empty((n, m), dtype=torch.float32, device="cuda").uniform_(-1, 1)
z = c.double()
v = torch.randn((m, r), dtype=torch.float64, device="cuda")
s = torch.empty((n, r), dtype=torch.float64, device="cuda")


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


w0 = torch.zeros((m, r), dtype=torch.float64, device="cuda")
w1 = torch.zeros((m, r), dtype=torch.float64, device="cuda")
ms64 = timed(lambda: ext.cg_matvec_accum(z, v, w0, s))
ms32 = timed(lambda: ext.cg_matvec_accum_c32(c, v, w1, 1.0, False, s))
w0.zero_(); w1.zero_()
ext.cg_matvec_accum(z, v, w0, s)
ext.cg_matvec_accum_c32(c, v, w1, 1.0, False, s)
err = float((w0 - w1).abs().max() / w0.abs().max())
flops = 4.0 * n * m * r
print(json.dumps({"n": n, "m": m, "r": r, "notail": bool(os.environ.get("XGPR_B200_MATVEC_NOTAIL")),
                  "f64_ms": ms64, "f64_tflops_useful": flops / ms64 / 1e9, "f64_GB_per_s": 2 * n * m * 8 / ms64 / 1e6,
                  "c32_ms": ms32, "c32_tflops_useful": flops / ms32 / 1e9, "c32_GB_per_s": 2 * n * m * 4 / ms32 / 1e6,
                  "rel_diff": err}))
