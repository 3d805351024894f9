"""One tensor-core Gram update (for ncu): python tools/prof_gram_tc.py [n] [m] [slices] [digit_bits]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from xgpr_b200 import ext

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
m = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
s = int(sys.argv[3]) if len(sys.argv) > 3 else 5
bits = int(sys.argv[4]) if len(sys.argv) > 4 else 8
torch.manual_seed(0)
z = torch.cos(torch.randn((n, m), device="cuda", dtype=torch.float64)) * (2.0 / m) ** 0.5
g = torch.zeros((m, m), dtype=torch.float64, device="cuda")
for _ in range(2):
    ext.ztz_accum_tc(z, g, num_slices=s, digit_bits=bits)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); ext.ztz_accum_tc(z, g, num_slices=s, digit_bits=bits); b.record(); torch.cuda.synchronize()
print("ms", a.elapsed_time(b))
