"""Per-step device time of one chunk of the srht preconditioner's accumulation pass (cfg2: 32,768 rows -> 16,384 RFFs, rank
512): featurise, Z^T y, fused SRHT sample, tensor-core A^T Z.   python tools/prof_precond_chunk.py [rows] [rffs]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from xgpr_b200 import ext
from xgpr_b200.kernels import RBF, SRHTCompressor

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
m = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
rank = 512
x = torch.randn((rows, 64), device="cuda", dtype=torch.float32)
y = torch.randn(rows, device="cuda", dtype=torch.float64)
kern = RBF((rows, 64), m, random_seed=123, device="cuda")
kern.set_hyperparams(np.asarray([np.log(np.sqrt(0.0767)), np.log(0.358 / 8 * 4.0)]), logspace=True)
comp = SRHTCompressor(rank, m, random_seed=123, device="cuda")


def timed(fn, reps=10):
    for _ in range(3):
        out = fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    return out, a.elapsed_time(b) / reps


res = {"rows": rows, "rffs": m}
z, res["featurise_ms"] = timed(lambda: kern.transform_x(x))
zty = torch.zeros(m, dtype=torch.float64, device="cuda")
_, res["zty_ms"] = timed(lambda: ext.zty_accum(z, y, zty))
cz, res["srht_sample_ms"] = timed(lambda: comp.transform_x(z))
acc = torch.zeros((rank, m), dtype=torch.float64, device="cuda")
_, res["atb_tc_ms"] = timed(lambda: ext.atb_accum_tc(cz, z, acc))
res["sum_ms"] = res["featurise_ms"] + res["zty_ms"] + res["srht_sample_ms"] + res["atb_tc_ms"]
# the same steps fed from the float32 encoding of Z (what the pass reads when the solve keeps the encoding resident)
c = torch.empty((rows, m), dtype=torch.float32, device="cuda")
_, res["encode_ms"] = timed(lambda: kern.resident_generate(x, c))
c[:, 0] = 1.0
scale = kern.feature_scale()
_, res["zty_c32_ms"] = timed(lambda: ext.zty_accum_c32(c, y, zty, scale, True))
cz2, res["srht_sample_c32_ms"] = timed(lambda: comp.transform_encoded(c, scale, True))
_, res["atb_tc_c32_ms"] = timed(lambda: ext.atb_accum_tc_c32(cz2, c, acc, scale, True))
res["sum_c32_ms"] = res["zty_c32_ms"] + res["srht_sample_c32_ms"] + res["atb_tc_c32_ms"]
res["chunk_GB"] = rows * m * 8 / 1e9
print(json.dumps(res))
