"""Second Nystrom pass acc += Z^T (Z Q): tensor-core route against the fp64 multi-RHS kernels.
python tools/prof_ztzq.py [n] [m] [r]"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from xgpr_b200 import ext

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
m = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
r = int(sys.argv[3]) if len(sys.argv) > 3 else 512
z = torch.randn((n, m), dtype=torch.float64, device="cuda") * m ** -0.5
q = torch.linalg.qr(torch.randn((m, r), dtype=torch.float64, device="cuda"))[0].contiguous()
acc = torch.zeros((m, r), dtype=torch.float64, device="cuda")
w = torch.zeros((m, r), dtype=torch.float64, device="cuda")
s = torch.empty((n, r), dtype=torch.float64, device="cuda")


def timed(fn, reps=3):
    fn(); fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


ms_tc = timed(lambda: ext.ztzq_accum_tc(z, q, acc))
ms_f64 = timed(lambda: ext.cg_matvec_accum(z, q, w, s))
acc.zero_(); ext.ztzq_accum_tc(z, q, acc)
w.zero_(); ext.cg_matvec_accum(z, q, w, s)
ref = z.T @ (z @ q)
print(json.dumps({"n": n, "m": m, "r": r, "ms_tc": ms_tc, "ms_f64": ms_f64,
                  "tflops_equiv_tc": 4.0 * n * m * r / ms_tc / 1e9,
                  "err_tc": float((acc - ref).abs().max() / ref.abs().max()),
                  "err_f64": float((w - ref).abs().max() / ref.abs().max())}))
