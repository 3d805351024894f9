"""Generates tests/golden/l1_*.npz from the REFERENCE's own CPU extension.

Run in the build container only (needs oracle/_ref, built by ``make -C oracle ref`` from the
sources under /root/reference):

    python tests/golden/make_golden.py

Every fixture stores the seeded inputs and the reference extension's outputs for one native
entry point of SURVEY.md §2.2 (cpu_rf_gen/xgpr_cpu_rfgen_cpp_ext.cpp:23-147).  The fixtures
travel with the repo; /root/reference does not.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import xgpr_cpu_rfgen_cpp_ext as ref  # noqa: E402

RADEM = np.asarray([-1, 1], dtype=np.int8)


def pw(d):
    return int(2 ** np.ceil(np.log2(max(d, 2))))


def draw(rng, p, f):
    reps = max(1, -(-f // p))
    radem = rng.choice(RADEM, size=(3, 1, reps * p), replace=True)
    chi = np.sqrt(rng.chisquare(p, size=f))
    return radem, chi


def save(name, **arrs):
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrs)
    print(name, {k: (v.shape, str(v.dtype)) for k, v in arrs.items() if hasattr(v, "shape")})


def main():
    rng = np.random.default_rng(123)

    # ---- FWHT 2-D and SRHT (transform_functions.cpp:58-80, :94-121)
    for tag, dt in (("f32", np.float32), ("f64", np.float64)):
        for p in (2, 8, 64, 256, 2048):
            x = rng.standard_normal((5, p)).astype(dt)
            y = x.copy()
            ref.cpuFastHadamardTransform2D(y)
            radem = rng.choice(RADEM, size=(p,), replace=True)
            z = x.copy()
            ref.cpuSRHT(z, radem)
            save(f"l1_fht_{tag}_{p}", x=x, fht=y, radem=radem, srht=z)

    # ---- RBF feature gen + gradient (rbf_ops.cpp:26-106, :135-221)
    cases = [(7, 20, 48), (5, 64, 160), (4, 130, 100), (3, 1, 6), (3, 513, 40), (2, 1500, 1100)]
    for tag, dt in (("f32", np.float32), ("f64", np.float64)):
        for (n, d, f) in cases:
            p = pw(d)
            x = rng.standard_normal((n, d)).astype(dt)
            radem, chi = draw(rng, p, f)
            chi = chi.astype(dt)
            for icpt in (True, False):
                out = np.zeros((n, 2 * f))
                ref.cpuRBFFeatureGen(x, out, radem, chi, icpt)
                gout = np.zeros((n, 2 * f))
                grad = np.zeros((n, 2 * f, 1))
                sigma = 0.8173
                ref.cpuRBFGrad(x, gout, grad, radem, chi, sigma, icpt)
                save(f"l1_rbf_{tag}_{n}x{d}_{f}_{int(icpt)}", x=x, radem=radem, chi=chi,
                     fit_intercept=np.asarray(icpt), out=out, sigma=np.asarray(sigma),
                     gout=gout, grad=grad)

    # ---- conv sum-pool, gradient, max-pool (rbf_convolution.cpp:22-139,:159-282; conv1d_operations.cpp:22-124)
    ccases = [(6, 12, 5, 3, 40), (4, 20, 21, 9, 300), (5, 9, 36, 1, 100), (3, 30, 4, 7, 16)]
    for tag, dt in (("f32", np.float32), ("f64", np.float64)):
        for (n, m, c, w, f) in ccases:
            p = pw(w * c)
            x = rng.standard_normal((n, m, c)).astype(dt)
            radem, chi = draw(rng, p, f)
            chi = chi.astype(dt)
            seqlen = rng.integers(w, m + 1, size=n).astype(np.int32)
            seqlen[0] = m
            seqlen[-1] = w
            arrs = dict(x=x, radem=radem, chi=chi, seqlen=seqlen, conv_width=np.asarray(w))
            for st in (0, 1, 2):
                out = np.zeros((n, 2 * f))
                ref.cpuConv1dFGen(x, out, radem, chi, seqlen, w, st)
                arrs[f"out{st}"] = out
                gout = np.zeros((n, 2 * f))
                grad = np.zeros((n, 2 * f, 1))
                ref.cpuConvGrad(x, gout, radem, chi, seqlen, grad, 1.37, w, st)
                arrs[f"gout{st}"] = gout
                arrs[f"grad{st}"] = grad
            arrs["sigma"] = np.asarray(1.37)
            # max-pool: F == num_rffs, R == reps*P exactly
            fm = 2 * f
            repsm = -(-fm // p)
            radem_m = rng.choice(RADEM, size=(3, 1, repsm * p), replace=True)
            chi_m = np.sqrt(rng.chisquare(p, size=fm)).astype(dt)
            mp = np.zeros((n, fm), np.float32)
            ref.cpuConv1dMaxpool(x, mp, radem_m, chi_m, seqlen, w)
            arrs.update(radem_m=radem_m, chi_m=chi_m, maxpool=mp)
            save(f"l1_conv_{tag}_{n}x{m}x{c}_w{w}_{f}", **arrs)

    # ---- MiniARD gradient (ard_ops.cpp:39-125)
    for tag, dt in (("f32", np.float32), ("f64", np.float64)):
        for (n, d, f, splits) in ((6, 10, 24, [4]), (5, 37, 50, [5, 20]), (4, 128, 64, [])):
            x = rng.standard_normal((n, d)).astype(dt)
            w = rng.standard_normal((f, d)).astype(dt)
            bounds = [0] + splits + [d]
            L = len(bounds) - 1
            smap = np.zeros(d, np.int32)
            svals = np.zeros(d, np.float64)
            hp = rng.uniform(0.2, 1.5, size=L)
            for l in range(L):
                smap[bounds[l]:bounds[l + 1]] = l
                svals[bounds[l]:bounds[l + 1]] = hp[l]
            for icpt in (True, False):
                out = np.zeros((n, 2 * f))
                grad = np.zeros((n, 2 * f, L))
                ref.cpuMiniARDGrad(x, out, w, smap, svals, grad, icpt)
                save(f"l1_ard_{tag}_{n}x{d}_{f}_L{L}_{int(icpt)}", x=x, w=w, smap=smap, svals=svals,
                     fit_intercept=np.asarray(icpt), out=out, grad=grad)


if __name__ == "__main__":
    main()
