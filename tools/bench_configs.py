#!/usr/bin/env python
"""Per-config kernel throughput on one B200 (device-resident inputs, CUDA events, best of `reps`).
Not the headline bench (bench.py): this is the per-row-of-SURVEY-§8 measurement table for DESIGN.md."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from xgpr_b200 import ext  # noqa: E402
from xgpr_b200.kernels import RBF, Conv1dRBF, GraphRBF, MiniARD, SRHTCompressor  # noqa: E402

dev = torch.device("cuda", 0)


def timeit(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


ONLY = os.environ.get("XB_CASE")          # substring filter on the case name (for ncu runs)


def report(name, rows, ms, bytes_per_row=None, flops=None, extra=None):
    d = {"case": name, "rows": rows, "ms": round(ms, 4), "rows_per_s": rows / (ms * 1e-3)}
    if bytes_per_row:
        d["GB_per_s"] = bytes_per_row * rows / (ms * 1e-3) / 1e9
    if flops:
        d["TFLOP_per_s"] = flops / (ms * 1e-3) / 1e12
    if extra:
        d.update(extra)
    print(json.dumps(d), flush=True)


def rbf_case(name, n, d, m, grad=False):
    if ONLY and ONLY not in name:
        return
    k = RBF((n, d), m, device="cuda")
    x = torch.randn((n, d), device=dev, dtype=torch.float32) * (0.5 / np.sqrt(d))
    out = torch.empty((n, m), dtype=torch.float64, device=dev)
    if grad:
        g = torch.empty((n, m, 1), dtype=torch.float64, device=dev)
        ms = timeit(lambda: ext.cudaRBFGrad(x, out, g, k.radem_diag, k.chi_arr, 0.5, True))
        report(name, n, ms, d * 4 + 2 * m * 8)
    else:
        ms = timeit(lambda: ext.cudaRBFFeatureGen(x, out, k.radem_diag, k.chi_arr, True))
        report(name, n, ms, d * 4 + m * 8)


def conv_case(name, cls, n, mlen, c, w, m, seq_lo, seq_hi, onehot=False, grad=False):
    if ONLY and ONLY not in name:
        return
    parms = {"conv_width": w} if cls is Conv1dRBF else {}
    k = cls((n, mlen, c), m, device="cuda", kernel_spec_parms=parms)
    rng = np.random.default_rng(123)
    if onehot:
        idx = torch.from_numpy(rng.integers(0, c, (n, mlen))).to(dev)
        x = torch.nn.functional.one_hot(idx, c).to(torch.float32).contiguous() * 0.3
    else:
        x = torch.rand((n, mlen, c), device=dev, dtype=torch.float32) * 0.3
    seqlen = rng.integers(seq_lo, seq_hi + 1, n).astype(np.int32)
    out = torch.zeros((n, m), dtype=torch.float64, device=dev)
    wins = int((seqlen - k.conv_width + 1).sum())
    if grad:
        g = torch.zeros((n, m, 1), dtype=torch.float64, device=dev)
        ms = timeit(lambda: ext.cudaConvGrad(x, out, k.radem_diag, k.chi_arr, seqlen, g, 0.5, k.conv_width, 0), reps=3, warm=1)
    else:
        ms = timeit(lambda: ext.cudaConv1dFGen(x, out, k.radem_diag, k.chi_arr, seqlen, k.conv_width, 0), reps=3, warm=1)
    report(name, n, ms, mlen * c * 4 + 4 + m * 8 * (2 if grad else 1),
           extra={"windows": wins, "freq_windows_per_s": wins * (m // 2) / (ms * 1e-3)})


def ard_case(n=65536, d=128, m=8192):
    """BASELINE config 5 (ii): MiniARD gradient features, 128 inputs -> 8,192 RFFs, two lengthscales; kernel only.
    Algorithmic bytes per row = 4 D + 8 M (features) + 8 M L (gradient); flops = 2 D F for the dense W x product."""
    name = f"cfg5 MiniARD grad {n}x{d} -> {m}, L=2 (kernel only)"
    if ONLY and ONLY not in name:
        return
    k = MiniARD((n, d), m, device="cuda", kernel_spec_parms={"split_points": [64]})
    k.set_hyperparams(np.log(np.asarray([0.277, 0.1, 0.07])), logspace=True)
    k.precompute_weights()
    x = torch.randn((n, d), dtype=torch.float32, device=dev)
    out = torch.zeros((n, m), dtype=torch.float64, device=dev)
    grad = torch.zeros((n, m, 2), dtype=torch.float64, device=dev)
    ms = timeit(lambda: ext.cudaMiniARDGrad(x, out, k.precomputed_weights, k.ard_position_key, k.full_ard_weights,
                                            grad, True), reps=3, warm=1)
    report(name, n, ms, d * 4 + m * 8 * 3, flops=2.0 * n * d * (m // 2))
    ms = timeit(lambda: ext.cudaMiniARDGrad(x, out, k.precomputed_weights, k.ard_position_key, k.full_ard_weights,
                                            grad, True, overwrite=True), reps=3, warm=1)
    report(name.replace("(kernel only)", "(overwrite form, kernel only)"), n, ms, d * 4 + m * 8 * 3,
           flops=2.0 * n * d * (m // 2))


def main():
    which = sys.argv[1:] or ["rbf", "conv", "gram", "misc"]
    if "rbf" in which:
        rbf_case("cfg1 RBF 20000x512 -> 2048", 20000, 512, 2048)
        rbf_case("cfg2 RBF 32768x64 -> 16384", 32768, 64, 16384)
        rbf_case("cfg5 RBF/MiniARD fwd 16384x128 -> 32768", 16384, 128, 32768)
        rbf_case("RBF grad 32768x64 -> 8192", 32768, 64, 8192, grad=True)
        rbf_case("RBF 8192x1024 -> 4096", 8192, 1024, 4096)
        rbf_case("RBF 4096x3001 -> 8192 (cta path)", 4096, 3001, 8192)
        rbf_case("big cfg1-shape RBF 262144x512 -> 2048", 262144, 512, 2048)
        rbf_case("big RBF 131072x1024 -> 4096", 131072, 1024, 4096)
        rbf_case("big RBF 262144x256 -> 2048", 262144, 256, 2048)
        rbf_case("big cta RBF 16384x3001 -> 8192 (cta path)", 16384, 3001, 8192)
    if "conv" in which:
        conv_case("cfg3 Conv1dRBF 2048x500x21 w9 -> 8192 (full length)", Conv1dRBF, 2048, 500, 21, 9, 8192, 500, 500, onehot=True)
        conv_case("cfg3 Conv1dRBF 2048x500x21 w9 -> 8192 (len 50..500)", Conv1dRBF, 2048, 500, 21, 9, 8192, 50, 500, onehot=True)
        conv_case("headline Conv1dRBF 2048x500x21 w9 -> 16384", Conv1dRBF, 2048, 500, 21, 9, 16384, 500, 500, onehot=True)
        conv_case("cfg4 GraphRBF 16384x64x36 -> 8192 (4..64 nodes)", GraphRBF, 16384, 64, 36, 1, 8192, 4, 64)
        conv_case("cfg4 GraphRBF 16384x64x36 -> 8192 (64 nodes)", GraphRBF, 16384, 64, 36, 1, 8192, 64, 64)
        conv_case("Conv1dRBF grad 1024x500x21 w9 -> 8192", Conv1dRBF, 1024, 500, 21, 9, 8192, 500, 500, onehot=True, grad=True)
    if "gram" in which:
        for n, m in [(4096, 2048), (4096, 8192)]:
            z = torch.randn((n, m), dtype=torch.float64, device=dev)
            c = torch.zeros((m, m), dtype=torch.float64, device=dev)
            ms = timeit(lambda: ext.ztz_accum(z, c), reps=3, warm=1)
            report(f"ZtZ {n}x{m} f64 (ours)", n, ms, flops=2.0 * n * m * m)
            ms = timeit(lambda: torch.addmm(c, z.T, z, out=c), reps=3, warm=1)
            report(f"ZtZ {n}x{m} f64 (torch/cuBLAS dgemm, comparator)", n, ms, flops=2.0 * n * m * m)
        n, m = 32768, 16384
        z = torch.randn((n, m), dtype=torch.float64, device=dev)
        y = torch.randn(n, dtype=torch.float64, device=dev)
        zty = torch.zeros(m, dtype=torch.float64, device=dev)
        ms = timeit(lambda: ext.zty_accum(z, y, zty))
        report(f"Zty {n}x{m}", n, ms, m * 8)
        for r in (1, 26):
            v = torch.randn((m, r), dtype=torch.float64, device=dev)
            w = torch.zeros((m, r), dtype=torch.float64, device=dev)
            scratch = torch.empty((n, r), dtype=torch.float64, device=dev)
            ms = timeit(lambda: ext.cg_matvec_accum(z, v, w, scratch))
            report(f"Zt(Zv) {n}x{m} r={r} (Z read twice)", n, ms, 2 * m * 8, flops=4.0 * n * m * r)
    if "misc" in which:
        n, p = 32768, 16384
        z = torch.randn((n, p), dtype=torch.float64, device=dev)
        comp = SRHTCompressor(512, p, device="cuda", double_precision=True)
        ms = timeit(lambda: ext.cudaSRHT(z, comp.radem))
        report(f"SRHT f64 {n}x{p} in place", n, ms, 2 * p * 8)
        x = torch.randn((1 << 20, 64), dtype=torch.float32, device=dev)
        ms = timeit(lambda: ext.cudaFastHadamardTransform2D(x))
        report("FHT f32 1Mx64 in place", 1 << 20, ms, 2 * 64 * 4)
        ard_case()
        k = MiniARD((4096, 128), 8192, device="cuda", kernel_spec_parms={"split_points": [64]})
        xa = torch.randn((4096, 128), dtype=torch.float32, device=dev)
        ms = timeit(lambda: k.gradient_x(xa), reps=3, warm=1)
        report("MiniARD grad 4096x128 -> 8192, L=2 (incl. allocs)", 4096, ms, 128 * 4 + 8192 * 8 * 3)


if __name__ == "__main__":
    main()
