"""Fused SRHT sample on an fp64 feature chunk (the compressor of the srht preconditioner): python tools/prof_srht_sample.py [n] [m] [rank]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from xgpr_b200.kernels import SRHTCompressor

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
m = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
rank = int(sys.argv[3]) if len(sys.argv) > 3 else 512
z = torch.randn((n, m), dtype=torch.float64, device="cuda")
comp = SRHTCompressor(rank, m, random_seed=123, device="cuda")
for _ in range(3):
    comp.transform_x(z)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    comp.transform_x(z)
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / 10
print(json.dumps({"n": n, "m": m, "rank": rank, "ms": ms, "GB_per_s_in": n * m * 8 / ms / 1e6}))
