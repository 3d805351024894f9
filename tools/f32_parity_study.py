#!/usr/bin/env python
"""float32 parity study (VERDICT r01 "weak" #1): how far are production-precision (float32 FHT) fitted weights and
predictions from the REFERENCE's, and how much of that distance is the MUFU cos/sin rather than unavoidable
glibc-vs-CUDA libm differences amplified by the conditioning of the solve?

Runs cfg1 (full size, exact fit), cfg2 (20k-row subsample, 16,384 RFFs, preconditioned CG) and the small l3 float32
fixture against the reference-generated fixtures at the three cos/sin levels of include/xgpr_b200.h (0 = Cody-Waite + MUFU,
1 = polynomial on the FMA pipe, 2 = CUDA sincosf), and times cfg2-shaped feature generation at each level.

    python tools/f32_parity_study.py > profiles/r02_f32_parity.json
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from config_inputs import tabular  # noqa: E402
from xgpr_b200 import ext, fitting  # noqa: E402
from xgpr_b200.kernels import RBF  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def rel(a, b):
    return float(np.abs(a - b).max() / np.abs(b).max())


def feature_error(kern, x, sigma):
    """max and rms |z - z_ref| / scale against the reference's CPU extension (oracle/_ref), when it is present."""
    sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
    try:
        import xgpr_cpu_rfgen_cpp_ext as ref
    except ImportError:
        return None
    host = RBF(x.shape, kern.get_num_rffs(), random_seed=123, device="cpu")
    xs = x.copy()
    xs *= np.float64(sigma)
    zr = np.zeros((x.shape[0], kern.get_num_rffs()))
    ref.cpuRBFFeatureGen(xs, zr, host.radem_diag, host.chi_arr, True)
    zr[:, 0] = 1.0
    z = kern.transform_x(x).cpu().numpy()
    scale = float(np.abs(zr[:, 1:]).max())
    d = (z - zr)[:, 1:] / scale
    return {"max": float(np.abs(d).max()), "rms": float(np.sqrt((d ** 2).mean()))}


def cfg1():
    g = np.load(os.path.join(GOLDEN, "cfg_1_exact.npz"))
    x, y = tabular(20000, 512)
    kern = RBF(x.shape, 2048, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(g["hyperparams"]), logspace=True)
    ds = fitting.build_shard_dataset(x, y, None, chunk_size=2000)
    w, _, _ = fitting.calc_weights_exact(ds, kern)
    pred = fitting.predict(kern, w, x[:500], None, 2000, ds.get_ymean(), ds.get_ystd())
    # conditioning of the solved system, from the Gram the GPU built
    ztz, _, _ = fitting.calc_design_mat(ds, kern)
    ev = torch.linalg.eigvalsh(ztz)
    lam2 = float(g["hyperparams"][0]) ** 2
    return {"weights_rel_max": rel(w.cpu().numpy(), g["weights"]),
            "pred_abs_max_over_ystd": float(np.abs(pred - g["pred"]).max()) / ds.get_ystd(),
            "cond_ztz_plus_2lam2": float((ev[-1] + 2 * lam2) / (max(float(ev[0]), 0.0) + 2 * lam2)),
            "features": feature_error(kern, x[:2000], float(g["hyperparams"][1]))}


def cfg2():
    g = np.load(os.path.join(GOLDEN, "cfg_2_cg.npz"))
    x, y = tabular(20000, 64)
    kern = RBF(x.shape, 16384, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(g["hyperparams"]), logspace=True)
    ds = fitting.build_shard_dataset(x, y, None, chunk_size=2000)
    pre = fitting.RandNysPreconditioner(kern, ds, 512, False, 123, "srht")
    w, nit, _ = fitting.cg_fit_lib_internal(kern, ds, 1e-6, 500, pre, False)
    xt, _ = tabular(200, 64, seed=321)
    pred = fitting.predict(kern, w, xt, None, 2000, ds.get_ymean(), ds.get_ystd())
    # the same system solved to 1e-10: separates the CG stopping error from the feature error
    w10, nit10, _ = fitting.cg_fit_lib_internal(kern, ds, 1e-10, 500, pre, False)
    return {"weights_rel_max": rel(w.cpu().numpy(), g["weights"]), "cg_iters": int(nit), "ref_cg_iters": int(g["niter"]),
            "ratio": float(pre.achieved_ratio), "ref_ratio": float(g["ratio"]),
            "pred_abs_max_over_ystd": float(np.abs(pred - g["pred"]).max()) / ds.get_ystd(),
            "weights_rel_max_tol1e-10_vs_ref_tol1e-6": rel(w10.cpu().numpy(), g["weights"]),
            "weights_rel_max_own_tol1e-6_vs_tol1e-10": rel(w.cpu().numpy(), w10.cpu().numpy()),
            "features": feature_error(kern, x[:1000], float(g["hyperparams"][1]))}


def l3_small():
    g = np.load(os.path.join(GOLDEN, "l3_fit_rbf_f32.npz"))
    kern = RBF(g["x"].shape, 256, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(g["hyperparams"]), logspace=True)
    ds = fitting.build_shard_dataset(g["x"], g["y"], None, chunk_size=97)
    w, _, _ = fitting.calc_weights_exact(ds, kern)
    return {"weights_rel_max": rel(w.cpu().numpy(), g["w_exact"])}


def speed(rows=262144, reps=5):
    kern = RBF((rows, 64), 16384, random_seed=123, device="cuda")
    g = torch.Generator(device="cuda")
    g.manual_seed(1)
    x = torch.randn((rows, 64), device="cuda", generator=g) * 0.2
    out = torch.empty((rows, 16384), dtype=torch.float64, device="cuda")
    for _ in range(2):
        ext.cudaRBFFeatureGen(x, out, kern.radem_diag, kern.chi_arr, True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        ext.cudaRBFFeatureGen(x, out, kern.radem_diag, kern.chi_arr, True)
    b.record()
    torch.cuda.synchronize()
    return {"rows_per_s": rows * reps / (a.elapsed_time(b) * 1e-3)}


def main():
    res = {}
    for mode, on in (("mufu", 0), ("poly", 1), ("sincosf", 2)):
        ext.set_accurate_sincos(on)
        res[mode] = {"cfg1": cfg1(), "cfg2": cfg2(), "l3_rbf_f32": l3_small(), "cfg2_feature_gen": speed()}
    ext.set_accurate_sincos(None)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
