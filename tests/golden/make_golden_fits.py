"""Generates tests/golden/l3_fit_*.npz from the REFERENCE's own L4 toolkits run on its CPU path.

Build container only (imports /root/reference/src/xGPR with the CPU extension from oracle/_ref):

    python tests/golden/make_golden_fits.py

Pins the downstream algebra of SURVEY.md Appendix A.8 for the mirror in xgpr_b200/fitting.py: exact-fit
weights (lambda^2 added twice), Nystrom preconditioner ("srht" and "srht_2") achieved ratio, CG-fit weights,
exact variance matrix and predictions (mean + variance), for float64 and float32 FHT.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import xgpr_cpu_rfgen_cpp_ext as _ext  # noqa: E402

sys.modules["xGPR.xgpr_cpu_rfgen_cpp_ext"] = _ext
sys.path.insert(0, "/root/reference/src")
from xGPR.kernels import KERNEL_NAME_TO_CLASS  # noqa: E402
from xGPR.data_handling.dataset_builder import build_regression_dataset  # noqa: E402
from xGPR.fitting_toolkit.exact_fitting_toolkit import calc_weights_exact, calc_variance_exact  # noqa: E402
from xGPR.fitting_toolkit.cg_fitting_toolkit import cg_fit_lib_internal  # noqa: E402
from xGPR.preconditioners.rand_nys_preconditioners import RandNysPreconditioner  # noqa: E402

CASES = [
    # name, kernel, xdim, num_rffs, parms, hyperparams [lambda, sigma], double_precision, rank, variance_rffs
    ("rbf_f64", "RBF", (600, 20), 256, {}, [0.5, 0.35], True, 32, 64),
    ("rbf_f32", "RBF", (600, 20), 256, {}, [0.5, 0.35], False, 32, 64),
    ("conv_f64", "Conv1dRBF", (240, 12, 5), 192, {"conv_width": 3}, [0.6, 0.5], True, 24, 32),
]


def main():
    for name, kname, xdim, num_rffs, parms, hp, dp, rank, vrffs in CASES:
        rng = np.random.default_rng(77)
        x = rng.standard_normal(xdim)
        seqlen = None
        if len(xdim) == 3:
            seqlen = rng.integers(parms["conv_width"], xdim[1] + 1, size=xdim[0]).astype(np.int32)
            y = np.sin(x[:, 0, :].sum(1)) + 0.1 * rng.standard_normal(xdim[0]) + 2.0
        else:
            y = np.sin(x[:, :4].sum(1)) + 0.1 * rng.standard_normal(xdim[0]) + 2.0
        xtest = rng.standard_normal((50,) + tuple(xdim[1:]))
        stest = None if seqlen is None else rng.integers(parms["conv_width"], xdim[1] + 1, size=50).astype(np.int32)
        dataset = build_regression_dataset(x, y, sequence_lengths=seqlen, chunk_size=97)
        kern = KERNEL_NAME_TO_CLASS[kname](xdim, num_rffs, random_seed=123, device="cpu", double_precision=dp,
                                           kernel_spec_parms=parms)
        kern.set_hyperparams(np.log(np.asarray(hp)), logspace=True)
        w_exact, _, _ = calc_weights_exact(dataset, kern)
        out = dict(x=x, y=y, xtest=xtest, hyperparams=np.asarray(hp), w_exact=w_exact,
                   trainy_mean=np.asarray(dataset.get_ymean()), trainy_std=np.asarray(dataset.get_ystd()))
        if seqlen is not None:
            out["seqlen"], out["stest"] = seqlen, stest
        for method in ("srht", "srht_2"):
            pre = RandNysPreconditioner(kern, dataset, rank, False, 123, method)
            w_cg, n_iter, _ = cg_fit_lib_internal(kern, dataset, cg_tol=1e-12, max_iter=500, preconditioner=pre,
                                                  verbose=False)
            out[f"ratio_{method}"] = np.asarray(pre.achieved_ratio)
            out[f"w_cg_{method}"] = w_cg
            out[f"niter_{method}"] = np.asarray(n_iter)
            out[f"zty_{method}"] = pre.z_trans_y
        w_cg0, n0, _ = cg_fit_lib_internal(kern, dataset, cg_tol=1e-12, max_iter=2000, preconditioner=None, verbose=False)
        out["w_cg_none"], out["niter_none"] = w_cg0, np.asarray(n0)
        var = calc_variance_exact(kern, dataset, vrffs)
        out["var"] = var
        lam = kern.get_lambda()
        z = kern.transform_x(xtest, stest)
        # prediction with the CG weights (what fit(mode="cg") stores), xgp_regression.py:118-134
        out["pred_cg"] = (z * w_cg0[None, :]).sum(1) * dataset.get_ystd() + dataset.get_ymean()
        out["pred_exact"] = (z * w_exact[None, :]).sum(1) * dataset.get_ystd() + dataset.get_ymean()
        zv = z[:, :vrffs]
        pv = lam ** 2 + lam ** 2 * (zv * (var @ zv.T).T).sum(1)
        pv[pv < 0] = 0
        out["pred_var"] = pv * dataset.get_ystd() ** 2
        np.savez_compressed(os.path.join(HERE, f"l3_fit_{name}.npz"), **out)
        print(name, "exact |w|", float(np.abs(w_exact).max()), "cg iters", {m: int(out[f"niter_{m}"]) for m in
              ("srht", "srht_2", "none")}, "ratios", float(out["ratio_srht"]), float(out["ratio_srht_2"]),
              "max|w_cg - w?|", float(np.abs(out["w_cg_srht"] - w_cg0).max()))


if __name__ == "__main__":
    main()
