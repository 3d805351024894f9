"""Generates tests/golden/l3_nmll_*.npz from the REFERENCE's own model class on its CPU path (build container only;
imports /root/reference/src/xGPR with the CPU extension from oracle/_ref):

    python tests/golden/make_golden_nmll.py

Pins the NMLL call sites mirrored in xgpr_b200/scoring.py: exact_nmll, exact_nmll_gradient (value + gradient w/r/t the
log hyperparameters, incl. a MiniARD kernel with two lengthscales) and approximate_nmll (SLQ, manual settings so that
the preconditioner rank fits the small problem).  Inputs are those of l3_fit_*.npz where a case of that name exists.
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
from xGPR import xGPRegression  # noqa: E402
from xGPR.kernels import KERNEL_NAME_TO_CLASS  # noqa: E402
from xGPR.data_handling.dataset_builder import build_regression_dataset  # noqa: E402

CASES = [
    # name, kernel, xdim, num_rffs, parms, hyperparams (not log), double_precision, approx settings
    ("rbf_f64", "RBF", (600, 20), 256, {}, [0.5, 0.35], True, {"max_rank": 32, "nsamples": 25, "nmll_tol": 1e-7}),
    ("rbf_f32", "RBF", (600, 20), 256, {}, [0.5, 0.35], False, {"max_rank": 32, "nsamples": 25, "nmll_tol": 1e-7}),
    ("conv_f64", "Conv1dRBF", (240, 12, 5), 192, {"conv_width": 3}, [0.6, 0.5], True,
     {"max_rank": 24, "nsamples": 25, "nmll_tol": 1e-7}),
    ("miniard_f64", "MiniARD", (500, 20), 128, {"split_points": [8]}, [0.5, 0.3, 0.45], True,
     {"max_rank": 32, "nsamples": 25, "nmll_tol": 1e-7}),
]


def main():
    for name, kname, xdim, num_rffs, parms, hp, dp, approx in CASES:
        fit = os.path.join(HERE, f"l3_fit_{name}.npz")
        seqlen = None
        if os.path.exists(fit):
            g = np.load(fit)
            x, y = g["x"], g["y"]
            seqlen = g["seqlen"] if "seqlen" in g.files else None
        else:
            rng = np.random.default_rng(78)
            x = rng.standard_normal(xdim)
            y = np.sin(x[:, :4].sum(1)) + 0.5 * np.cos(x[:, 10:14].sum(1)) + 0.1 * rng.standard_normal(xdim[0]) + 2.0
        dataset = build_regression_dataset(x, y, sequence_lengths=seqlen, chunk_size=97)
        model = xGPRegression(num_rffs=num_rffs, variance_rffs=16, kernel_choice=kname, device="cpu",
                              kernel_settings=parms, verbose=False, random_seed=123)
        model.kernel = KERNEL_NAME_TO_CLASS[kname](x.shape, num_rffs, random_seed=123, device="cpu",
                                                   double_precision=dp, kernel_spec_parms=parms)
        loghp = np.log(np.asarray(hp))
        out = dict(x=x, y=y, hyperparams=np.asarray(hp))
        if seqlen is not None:
            out["seqlen"] = seqlen
        out["nmll_exact"] = np.asarray(model.exact_nmll(loghp.copy(), dataset))
        val, grad = model.exact_nmll_gradient(loghp.copy(), dataset)
        out["nmll_grad_value"], out["nmll_grad"] = np.asarray(val), np.asarray(grad)
        if seqlen is not None:      # the reference's subsampling indexes sequence_lengths, so it needs sequence data
            val_s, grad_s = model.exact_nmll_gradient(loghp.copy(), dataset, subsample=0.5)
            out["nmll_grad_value_sub"], out["nmll_grad_sub"] = np.asarray(val_s), np.asarray(grad_s)
        settings = {"max_rank": approx["max_rank"], "nsamples": approx["nsamples"], "nmll_tol": approx["nmll_tol"],
                    "nmll_iter": 500, "preconditioner_mode": "srht_2"}
        out["nmll_approx"] = np.asarray(model.approximate_nmll(loghp.copy(), dataset, manual_settings=dict(settings)))
        out["approx_max_rank"], out["approx_nsamples"], out["approx_tol"] = (np.asarray(approx["max_rank"]),
                                                                              np.asarray(approx["nsamples"]),
                                                                              np.asarray(approx["nmll_tol"]))
        np.savez_compressed(os.path.join(HERE, f"l3_nmll_{name}.npz"), **out)
        print(name, "exact", float(out["nmll_exact"]), "grad", out["nmll_grad"], "sub", out.get("nmll_grad_sub"),
              "approx", float(out["nmll_approx"]))


if __name__ == "__main__":
    main()
