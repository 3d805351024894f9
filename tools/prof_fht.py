"""FWHT / SRHT throughput at the CTA-kernel sizes: python tools/prof_fht.py"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from xgpr_b200 import ext

for dt, n, p in [(torch.float64, 32768, 16384), (torch.float64, 65536, 8192), (torch.float32, 65536, 16384),
                 (torch.float32, 262144, 4096), (torch.float64, 262144, 2048), (torch.float64, 16384, 32768)]:
    x = torch.randn((n, p), dtype=dt, device="cuda")
    radem = (torch.randint(0, 2, (p,), device="cuda", dtype=torch.int8) * 2 - 1)
    for name, fn in (("fht", lambda: ext.cudaFastHadamardTransform2D(x)), ("srht", lambda: ext.cudaSRHT(x, radem))):
        fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        print(json.dumps({"op": name, "dtype": str(dt), "n": n, "p": p, "ms": ms,
                          "GB_per_s": 2 * x.numel() * x.element_size() / ms / 1e6}))
    del x
