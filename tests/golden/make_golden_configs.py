"""Generates tests/golden/cfg_*.npz: the BASELINE.json configurations at their REAL widths, run through the REFERENCE's own
model class (xGPRegression, device="cpu") in the build container:

    python tests/golden/make_golden_configs.py [cfg1] [cfg2] [cfg5] [conv]

  cfg1   RBF 20,000 x 512 -> 2,048 RFFs, fit(mode="exact") + predict: weights, 500 predictions, 4 feature rows.
         FULL SIZE (BASELINE.json configs[0] is exactly this CPU run).
  cfg2   RBF 20,000-row subsample x 64 -> 16,384 RFFs: rank-512 "srht" preconditioner (achieved ratio), fit(mode="cg",
         tol=1e-6) weights + iteration count + 200 predictions, and _check_rank_ratio(0.2, 512) -- the reference caps the
         ratio pass at 8,192 RFFs (model_baseclass.py:438-475), which is what pins xgpr_b200.scoring.check_rank_ratio.
  cfg5   MiniARD (split_points=[64]) 20,000-row subsample x 128 -> 32,768 RFFs: approximate_nmll (SLQ, 25 probes, rank-512
         "srht_2", tol 1e-6).
  conv   cfg3 / cfg4 / the Conv1dRBF @16,384 headline at their real sequence shapes: a handful of ragged sequences through
         the reference's CPU extension (cpuConv1dFGen), all three scalings.  (tests/test_configs_gpu.py also compares
         against oracle/_ref LIVE on the GPU box; these fixtures keep the comparison alive where oracle/_ref is absent.)

Inputs are regenerated in the tests from the seeds below (NumPy Generator streams are platform independent); only the
reference's OUTPUTS are stored, so the fixtures stay small.
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import xgpr_cpu_rfgen_cpp_ext as _ext  # noqa: E402

sys.modules["xGPR.xgpr_cpu_rfgen_cpp_ext"] = _ext
sys.path.insert(0, "/root/reference/src")
from xGPR import xGPRegression  # noqa: E402
from xGPR.data_handling.dataset_builder import build_regression_dataset  # noqa: E402

sys.path.insert(0, os.path.join(ROOT, "tests"))
from config_inputs import CONV_CASES, LAMBDA, conv_cols, conv_inputs, tabular  # noqa: E402


def cfg1():
    x, y = tabular(20000, 512)
    hp = np.asarray([LAMBDA, 0.358 / np.sqrt(512) * 4.0])
    ds = build_regression_dataset(x, y, chunk_size=2000)
    model = xGPRegression(num_rffs=2048, variance_rffs=64, kernel_choice="RBF", device="cpu", verbose=False,
                          random_seed=123)
    model.set_hyperparams(np.log(hp), ds)
    t = time.time()
    model.fit(ds, mode="exact", suppress_var=True)
    pred = model.predict(x[:500])
    z = model.kernel.transform_x(x[:4])
    print("cfg1 exact fit %.1fs  |w|max %.4g" % (time.time() - t, np.abs(model.weights).max()))
    np.savez_compressed(os.path.join(HERE, "cfg_1_exact.npz"), hyperparams=hp, weights=model.weights, pred=pred,
                        z_first4=z, trainy_mean=np.asarray(ds.get_ymean()), trainy_std=np.asarray(ds.get_ystd()))


def cfg2():
    x, y = tabular(20000, 64)
    hp = np.asarray([LAMBDA, 0.358 / np.sqrt(64) * 4.0])
    ds = build_regression_dataset(x, y, chunk_size=2000)
    model = xGPRegression(num_rffs=16384, variance_rffs=64, kernel_choice="RBF", device="cpu", verbose=False,
                          random_seed=123)
    model.set_hyperparams(np.log(hp), ds)
    t = time.time()
    pre, ratio = model.build_preconditioner(ds, max_rank=512, method="srht")
    n_iter, losses = model.fit(ds, preconditioner=pre, mode="cg", tol=1e-6, suppress_var=True, run_diagnostics=True)
    weights = model.weights.copy()
    xt, _ = tabular(200, 64, seed=321)
    pred = model.predict(xt)
    print("cfg2 cg fit %.1fs  ratio %.4g  iters %d  |w|max %.4g" % (time.time() - t, ratio, n_iter, np.abs(weights).max()))
    t = time.time()
    rr = model._check_rank_ratio(ds, 0.2, 512)          # runs on min(num_rffs, 8192) RFFs
    print("cfg2 rank-ratio pass %.1fs  ratio %.6g  (kernel restored to %d RFFs)" % (time.time() - t, rr,
                                                                                   model.kernel.get_num_rffs()))
    np.savez_compressed(os.path.join(HERE, "cfg_2_cg.npz"), hyperparams=hp, weights=weights, niter=np.asarray(n_iter),
                        ratio=np.asarray(ratio), pred=pred, losses=np.asarray(losses), rank_ratio_512=np.asarray(rr),
                        trainy_mean=np.asarray(ds.get_ymean()), trainy_std=np.asarray(ds.get_ystd()))


def cfg5():
    x, y = tabular(20000, 128)
    s = 0.358 / np.sqrt(128) * 4.0
    hp = np.asarray([LAMBDA, s, 0.7 * s])
    ds = build_regression_dataset(x, y, chunk_size=2000)
    model = xGPRegression(num_rffs=32768, variance_rffs=64, kernel_choice="MiniARD", device="cpu", verbose=False,
                          random_seed=123, kernel_settings={"split_points": [64]})
    model.set_hyperparams(np.log(hp), ds)
    settings = {"max_rank": 512, "nsamples": 25, "nmll_tol": 1e-6, "nmll_iter": 500, "preconditioner_mode": "srht_2"}
    t = time.time()
    val = model.approximate_nmll(np.log(hp), ds, manual_settings=dict(settings))
    print("cfg5 approximate_nmll %.1fs  value %.10g" % (time.time() - t, val))
    np.savez_compressed(os.path.join(HERE, "cfg_5_nmll.npz"), hyperparams=hp, nmll_approx=np.asarray(val),
                        max_rank=np.asarray(512), nsamples=np.asarray(25), tol=np.asarray(1e-6))


def conv():
    from xGPR.kernels import KERNEL_NAME_TO_CLASS
    for name, kind, n, w, m, sigma in CONV_CASES:
        x, lens = conv_inputs(kind, n, seed=31)
        cls = KERNEL_NAME_TO_CLASS["Conv1dRBF" if w > 1 else "GraphRBF"]
        out = {}
        for avg in ("none", "sqrt", "full"):
            parms = {"conv_width": w, "averaging": avg} if w > 1 else {"averaging": avg}
            k = cls(x.shape, m, random_seed=123, device="cpu", kernel_spec_parms=parms)
            k.set_hyperparams(np.log(np.asarray([LAMBDA, sigma])), logspace=True)
            t = time.time()
            z = k.transform_x(x, lens)
            cols = conv_cols(m)
            out["z_" + avg] = z[:, cols]              # a column subset keeps the fixture small; the GPU test also
                                                      # compares ALL columns against oracle/_ref live
            print(name, avg, "%.1fs" % (time.time() - t), z.shape, float(np.abs(z).max()))
        np.savez_compressed(os.path.join(HERE, f"cfg_conv_{name}.npz"), sigma=np.asarray(sigma), seqlen=lens, **out)


if __name__ == "__main__":
    which = sys.argv[1:] or ["cfg1", "cfg2", "cfg5", "conv"]
    for nm in which:
        {"cfg1": cfg1, "cfg2": cfg2, "cfg5": cfg5, "conv": conv}[nm]()
