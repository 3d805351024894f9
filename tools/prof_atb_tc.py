"""One Nystrom sketch accumulation acc (rank, M) += A^T Z on the tensor cores: python tools/prof_atb_tc.py [n] [m] [rank]"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from xgpr_b200 import ext

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
m = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
r = int(sys.argv[3]) if len(sys.argv) > 3 else 512
z = (torch.rand((n, m), dtype=torch.float64, device="cuda") * 2 - 1) * (2.0 / m) ** 0.5
a = torch.randn((n, r), dtype=torch.float64, device="cuda")
acc = torch.zeros((r, m), dtype=torch.float64, device="cuda")
for _ in range(2):
    ext.atb_accum_tc(a, z, acc)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    ext.atb_accum_tc(a, z, acc)
e1.record(); torch.cuda.synchronize()
print(json.dumps({"n": n, "m": m, "rank": r, "ms_per_call": e0.elapsed_time(e1) / 3}))
