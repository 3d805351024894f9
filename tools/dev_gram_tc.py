"""Development check + timing of the tcgen05 split-int8 Gram (xgpr_b200_ztz_accum_tc_f64) against fp64 references."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from xgpr_b200 import ext


def timed(fn, reps=3):
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def main():
    torch.manual_seed(0)
    dev = "cuda"
    # exact case: two-digit dyadic inputs
    for (n, m) in [(128, 256), (300, 200), (1000, 700), (4096, 1024)]:
        z = torch.randint(-8000, 8001, (n, m), device=dev, dtype=torch.float64) / 16384.0
        ref = z.T @ z
        for s in (3, 6):
            g = torch.zeros((m, m), dtype=torch.float64, device=dev)
            ext.ztz_accum_tc(z, g, num_slices=s)
            torch.cuda.synchronize()
            print(json.dumps({"case": "dyadic", "n": n, "m": m, "s": s, "max_abs_diff": float((g - ref).abs().max()),
                              "sym": float((g - g.T).abs().max())}))
    # feature-like inputs
    for (n, m) in [(2000, 2048), (5000, 1000)]:
        z = torch.cos(torch.randn((n, m), device=dev, dtype=torch.float64) * 3).float().double() * (2.0 / m) ** 0.5
        z[:, 0] = 1.0
        ref = z.T @ z
        for s, bits in ((4, 7), (5, 7), (6, 7), (7, 7), (4, 8), (5, 8), (6, 8)):
            g = torch.zeros((m, m), dtype=torch.float64, device=dev)
            ext.ztz_accum_tc(z, g, num_slices=s, digit_bits=bits)
            err = (g - ref).abs()
            d = ref.diagonal().sqrt()
            print(json.dumps({"case": "features", "n": n, "m": m, "s": s, "bits": bits, "max_abs": float(err.max()),
                              "max_rel_to_diag": float((err / (d[:, None] * d[None, :])).max())}))
    # timing
    for (n, m) in [(32768, 8192), (32768, 4096), (65536, 2048)]:
        z = torch.cos(torch.randn((n, m), device=dev, dtype=torch.float64)) * (2.0 / m) ** 0.5
        g = torch.zeros((m, m), dtype=torch.float64, device=dev)
        t64 = timed(lambda: ext.ztz_accum(z, g), 2)
        tt = timed(lambda: torch.addmm(g, z.T, z, out=g), 2)
        row = {"case": "timing", "n": n, "m": m, "dfma_ms": t64, "cublas_dgemm_ms": tt,
               "dfma_tflops": 2 * n * m * m / t64 / 1e9}
        for s, bits in ((4, 7), (6, 7), (5, 8), (4, 8)):
            ts = timed(lambda: ext.ztz_accum_tc(z, g, num_slices=s, digit_bits=bits), 3)
            row[f"tc_s{s}b{bits}_ms"] = ts
            row[f"tc_s{s}b{bits}_eff_tflops"] = 2 * n * m * m / ts / 1e9
        print(json.dumps(row))


if __name__ == "__main__":
    main()
