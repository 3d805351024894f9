"""Random-shape sweep of the consumers against NumPy with guard bands on every written array (the pool has no
compute-sanitizer):  python tools/fuzz_consumers.py [cases] [seed]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from xgpr_b200 import ext

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 200
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
PAD = 4096


def guarded(shape, fill=0.0):
    n = int(np.prod(shape))
    buf = torch.full((n + 2 * PAD,), float("nan"), dtype=torch.float64, device="cuda")
    v = buf[PAD:PAD + n].view(shape)
    v.fill_(fill)
    return buf, v, n


def intact(buf, n):
    return bool(torch.isnan(buf[:PAD]).all()) and bool(torch.isnan(buf[PAD + n:]).all())


worst = 0.0
for it in range(cases):
    n = int(rng.integers(1, 3000))
    m = 4 * int(rng.integers(1, 2500)) if rng.random() < 0.8 else int(rng.integers(1, 3000))
    r = int(rng.choice([1, 2, 3, 7, 8, 16, 24, 25, 26, 27, 31, 32, 33, 40, 58, 64, 70]))
    zf = rng.standard_normal((n, m)).astype(np.float32)
    zd = zf.astype(np.float64)
    v = rng.standard_normal((m, r))
    y = rng.standard_normal(n)
    want = zd.T @ (zd @ v)
    scale = float(np.abs(want).max()) + 1e-300
    z = torch.from_numpy(zd).cuda()
    vt = torch.from_numpy(v).cuda()
    wb, w, wn = guarded((m, r))
    sb, s, sn = guarded((n, r), 3.0)
    ext.cg_matvec_accum(z, vt, w, s)
    torch.cuda.synchronize()
    e1 = float(np.abs(w.cpu().numpy() - want).max()) / scale
    ok = intact(wb, wn) and intact(sb, sn)
    e2 = 0.0
    if m % 4 == 0:
        wb2, w2, _ = guarded((m, r))
        sb2, s2, _ = guarded((n, r), 3.0)
        ext.cg_matvec_accum_c32(torch.from_numpy(zf).cuda(), vt, w2, 1.0, False, s2)
        torch.cuda.synchronize()
        e2 = float(np.abs(w2.cpu().numpy() - want).max()) / scale
        ok = ok and intact(wb2, wn) and intact(sb2, sn)
    tb, t, tn = guarded((m,))
    ext.zty_accum(z, torch.from_numpy(y).cuda(), t)
    torch.cuda.synchronize()
    wy = zd.T @ y
    e3 = float(np.abs(t.cpu().numpy() - wy).max()) / (float(np.abs(wy).max()) + 1e-300)
    ok = ok and intact(tb, tn)
    worst = max(worst, e1, e2, e3)
    if not ok or max(e1, e2, e3) > 1e-11:
        print(json.dumps({"FAIL": [n, m, r], "f64": e1, "c32": e2, "zty": e3, "guards": ok}))
        sys.exit(1)
print(json.dumps({"cases": cases, "worst_rel_err": worst, "guards": "intact"}))
