"""Random-shape sweep of the encoded-input preconditioner steps (Z^T y, fused SRHT sample, tensor-core sketch) against their
fp64-input counterparts on Z = double(c) * scale, with guard bands on the written arrays:
    python tools/fuzz_precond_c32.py [cases] [seed]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from xgpr_b200 import ext
from xgpr_b200.kernels import SRHTCompressor

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 60
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
PAD = 4096


def guarded(shape):
    n = int(np.prod(shape))
    buf = torch.full((n + 2 * PAD,), float("nan"), dtype=torch.float64, device="cuda")
    v = buf[PAD:PAD + n].view(shape)
    v.zero_()
    return buf, v, n


def intact(buf, n):
    return bool(torch.isnan(buf[:PAD]).all()) and bool(torch.isnan(buf[PAD + n:]).all())


worst = {"zty": 0.0, "atb": 0.0}
for it in range(cases):
    n = int(rng.integers(1, 2500))
    m = 4 * int(rng.integers(130, 4200))                 # 520 .. 16,796 features
    icpt = bool(rng.integers(0, 2))
    scale = float(np.float32(1.0 / np.sqrt(m / 2 - 0.5)))
    cf = rng.uniform(-1, 1, (n, m)).astype(np.float32)
    if icpt:
        cf[:, 0] = 1.0
    c = torch.from_numpy(cf).cuda()
    s = np.full(m, scale)
    if icpt:
        s[0] = 1.0
    zd = cf.astype(np.float64) * s[None, :]
    z = torch.from_numpy(zd).cuda()
    y = torch.from_numpy(rng.standard_normal(n)).cuda()
    tb, t, tn = guarded((m,))
    ext.zty_accum_c32(c, y, t, scale, icpt)
    t64 = torch.zeros(m, dtype=torch.float64, device="cuda")
    ext.zty_accum(z, y, t64)
    torch.cuda.synchronize()
    e1 = float((t - t64).abs().max() / (t64.abs().max() + 1e-300))
    ok = intact(tb, tn)
    rank = int(rng.integers(2, min(600, m - 1)))
    comp = SRHTCompressor(rank, m, random_seed=int(rng.integers(1, 1000)), device="cuda")
    s64 = comp.transform_x(z)
    s32 = comp.transform_encoded(c, scale, icpt)
    same = True if s32 is None else bool(torch.equal(s64, s32))
    ab, a, an = guarded((rank, m))
    ext.atb_accum_tc_c32(s64, c, a, scale, icpt)
    a64 = torch.zeros((rank, m), dtype=torch.float64, device="cuda")
    ext.atb_accum_tc(s64, z, a64)
    torch.cuda.synchronize()
    want = s64.cpu().numpy().T @ zd
    e2 = float(np.abs(a.cpu().numpy() - want).max() / (np.abs(want).max() + 1e-300))
    e2b = float(np.abs(a64.cpu().numpy() - want).max() / (np.abs(want).max() + 1e-300))
    ok = ok and intact(ab, an)
    worst["zty"] = max(worst["zty"], e1)
    worst["atb"] = max(worst["atb"], e2)
    if not ok or not same or e1 > 1e-12 or e2 > 1e-9 or e2 > 50 * e2b + 1e-12:
        print(json.dumps({"FAIL": [n, m, rank, icpt], "zty": e1, "atb_c32": e2, "atb_f64": e2b, "srht_equal": same,
                          "guards": ok}))
        sys.exit(1)
print(json.dumps({"cases": cases, "worst": worst, "srht": "bit-identical", "guards": "intact"}))
