"""Generates tests/golden/l2_*.npz from the REFERENCE's Python kernel classes (CPU device).

Build container only: imports /root/reference/src/xGPR with its compiled CPU extension aliased in from
oracle/_ref.  Pins (a) the host parameter draws (radem_diag / chi_arr) and (b) transform_x / gradient_x
outputs of every registered kernel for the xgpr_b200.kernels mirror.

    python tests/golden/make_golden_kernels.py
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
from xGPR.kernels.srht_compressor import SRHTCompressor  # noqa: E402

CASES = [
    # name, kernel, xdim, num_rffs, parms, hyperparams
    ("rbf", "RBF", (11, 20), 96, {}, [0.3, 0.7]),
    ("rbf_nointercept", "RBF", (7, 84), 300, {"intercept": False}, [0.3, 0.2]),
    ("matern", "Matern", (9, 33), 128, {"matern_nu": 2.5}, [0.5, 0.4]),
    ("cauchy", "Cauchy", (9, 33), 128, {}, [0.5, 0.4]),
    ("miniard", "MiniARD", (10, 24), 160, {"split_points": [5, 13]}, [0.4, 0.5, 1.1, 0.8]),
    ("linear", "Linear", (6, 12), 64, {}, [0.7]),
    ("conv1drbf", "Conv1dRBF", (8, 15, 6), 200, {"conv_width": 4}, [0.4, 0.6]),
    ("conv1drbf_sqrt", "Conv1dRBF", (8, 15, 6), 200, {"conv_width": 4, "averaging": "sqrt"}, [0.4, 0.6]),
    ("conv1drbf_full", "Conv1dRBF", (8, 15, 6), 200, {"conv_width": 4, "averaging": "full"}, [0.4, 0.6]),
    ("conv1dmatern", "Conv1dMatern", (6, 12, 5), 100, {"conv_width": 3, "matern_nu": 1.5}, [0.4, 0.6]),
    ("conv1dcauchy", "Conv1dCauchy", (6, 12, 5), 100, {"conv_width": 3}, [0.4, 0.6]),
    ("graphrbf", "GraphRBF", (7, 10, 9), 140, {}, [0.4, 0.9]),
    ("graphmatern", "GraphMatern", (7, 10, 9), 140, {"matern_nu": 0.5}, [0.4, 0.9]),
    ("graphcauchy", "GraphCauchy", (7, 10, 9), 140, {}, [0.4, 0.9]),
    ("twolayer", "Conv1dTwoLayer", (6, 14, 5), 120, {"conv_width": 3, "init_rffs": 64}, [0.4, 0.3]),
]


def main():
    rng = np.random.default_rng(2024)
    for name, kname, xdim, num_rffs, parms, hp in CASES:
        kern = KERNEL_NAME_TO_CLASS[kname](xdim, num_rffs, random_seed=123, device="cpu",
                                           double_precision=False, kernel_spec_parms=parms)
        kern.set_hyperparams(np.log(np.asarray(hp)), logspace=True)
        x = rng.standard_normal(xdim)
        seqlen = None
        if len(xdim) == 3:
            w = parms.get("conv_width", 1)
            seqlen = rng.integers(w, xdim[1] + 1, size=xdim[0]).astype(np.int32)
            seqlen[0] = xdim[1]
        z = kern.transform_x(x, seqlen)
        zg, dz = kern.gradient_x(x, seqlen)
        arrs = dict(x=x, z=z, zg=zg, dz=dz, hyperparams=np.asarray(hp), num_rffs=np.asarray(num_rffs),
                    xdim=np.asarray(xdim))
        if seqlen is not None:
            arrs["seqlen"] = seqlen
        for attr in ("radem_diag", "chi_arr", "radem_diag1", "chi_arr1", "radem_diag2", "chi_arr2",
                     "precomputed_weights", "full_ard_weights", "ard_position_key"):
            v = getattr(kern, attr, None)
            if v is not None:
                arrs[attr] = np.asarray(v)
        np.savez_compressed(os.path.join(HERE, f"l2_{name}.npz"), **arrs)
        print(name, z.shape, dz.shape)

    comp = SRHTCompressor(24, 100, device="cpu", double_precision=True, random_seed=123)
    feats = rng.standard_normal((9, 100))
    np.savez_compressed(os.path.join(HERE, "l2_srht.npz"), x=feats, radem=comp.radem, col_sampler=comp.col_sampler,
                        out=comp.transform_x(feats), full=comp.transform_x(feats, no_compression=True))
    print("srht done")


if __name__ == "__main__":
    main()
