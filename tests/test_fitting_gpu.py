"""GPU parity of the downstream path: xgpr_b200/fitting.py with the CUDA
kernels and the C-ABI consumers of Z against fixtures produced by the REFERENCE's own toolkits on its CPU path
(tests/golden/make_golden_fits.py).

Tolerances: float64 FHT -> 1e-6 on fitted weights, variance matrix and predictions (north_star).  float32 FHT: the
features agree to ~5e-7 of scale, and the solve amplifies that by the conditioning of Z^T Z + lambda^2 I, so the
end-to-end float32 bound is looser and is stated where it is asserted; the float32 algebra on IDENTICAL features
is checked at 1e-6 separately (test_consumers_vs_numpy)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, load_golden

pytestmark = pytest.mark.gpu

CASES = {"rbf_f64": ("RBF", 256, {}, True, 32, 64), "rbf_f32": ("RBF", 256, {}, False, 32, 64),
         "conv_f64": ("Conv1dRBF", 192, {"conv_width": 3}, True, 24, 32)}


def _setup(name, resident=False):
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import KERNEL_NAME_TO_CLASS
    g = load_golden(os.path.join(GOLDEN, f"l3_fit_{name}.npz"))
    kname, m, parms, dp, rank_p, vrffs = CASES[name]
    kern = KERNEL_NAME_TO_CLASS[kname](g["x"].shape, m, random_seed=123, device="cuda", double_precision=dp,
                                       kernel_spec_parms=parms)
    kern.set_hyperparams(np.log(g["hyperparams"]), logspace=True)
    x = torch.from_numpy(g["x"]).cuda() if resident else g["x"]
    ds = fitting.build_shard_dataset(x, g["y"], g.get("seqlen"), chunk_size=97)
    return fitting, g, kern, ds, rank_p, vrffs


@pytest.mark.parametrize("name", sorted(CASES))
def test_fit_matches_reference(name):
    fitting, g, kern, ds, rank_p, vrffs = _setup(name)
    f64 = CASES[name][3]
    tol = dict(rtol=1e-6, atol=1e-9) if f64 else dict(rtol=2e-3, atol=2e-4)   # float32: see module docstring
    np.testing.assert_allclose(ds.get_ymean(), float(g["trainy_mean"]), rtol=1e-12)
    np.testing.assert_allclose(ds.get_ystd(), float(g["trainy_std"]), rtol=1e-12)
    w, _, _ = fitting.calc_weights_exact(ds, kern)
    np.testing.assert_allclose(w.cpu().numpy(), g["w_exact"], **tol)
    for method in ("srht", "srht_2"):
        pre = fitting.RandNysPreconditioner(kern, ds, rank_p, False, 123, method)
        np.testing.assert_allclose(pre.achieved_ratio, float(g[f"ratio_{method}"]), rtol=1e-6 if f64 else 1e-3)
        np.testing.assert_allclose(pre.get_zty().cpu().numpy(), g[f"zty_{method}"], rtol=1e-6 if f64 else 1e-3,
                                   atol=1e-9 if f64 else 1e-3)
        wcg, nit, _ = fitting.cg_fit_lib_internal(kern, ds, 1e-12, 500, pre, False)
        np.testing.assert_allclose(wcg.cpu().numpy(), g[f"w_cg_{method}"], **tol)
        assert abs(nit - int(g[f"niter_{method}"])) <= 2
    wcg0, nit0, _ = fitting.cg_fit_lib_internal(kern, ds, 1e-12, 2000, None, False)
    np.testing.assert_allclose(wcg0.cpu().numpy(), g["w_cg_none"], **tol)
    var = fitting.calc_variance_exact(kern, ds, vrffs)
    np.testing.assert_allclose(var.cpu().numpy(), g["var"], rtol=1e-6 if f64 else 1e-2, atol=1e-10 if f64 else 1e-5)
    pred, pvar = fitting.predict(kern, wcg0, g["xtest"], g.get("stest"), 23, ds.get_ymean(), ds.get_ystd(), var, vrffs)
    np.testing.assert_allclose(pred, g["pred_cg"], rtol=1e-6 if f64 else 1e-4, atol=1e-9 if f64 else 1e-4)
    np.testing.assert_allclose(pvar, g["pred_var"], rtol=1e-6 if f64 else 1e-3, atol=1e-10)
    pe = fitting.predict(kern, w, g["xtest"], g.get("stest"), 50, ds.get_ymean(), ds.get_ystd())
    np.testing.assert_allclose(pe, g["pred_exact"], rtol=1e-6 if f64 else 1e-4, atol=1e-9 if f64 else 1e-4)


@pytest.mark.parametrize("m", [256, 4096])
def test_device_cg_matches_host_loop(m):
    """The stream-ordered CG (scalar recurrences on the device, CUDA-graph replay, lagged convergence polling:
    ConjugateGrad._fit_device) against the host-driven loop that mirrors cg_tools.py:57-156 line by line: same iteration
    count, same loss history, same weights.  m = 256 runs the fused regenerate-per-iteration matvec, m = 4096 the resident
    float32 encoding with the gated cluster matvec."""
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(3)
    x = rng.standard_normal((3000, 20)).astype(np.float32)
    y = np.sin(x[:, :4].sum(1)) + 0.1 * rng.standard_normal(3000)
    kern = RBF(x.shape, m, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(np.asarray([0.5, 0.35])), logspace=True)
    ds = fitting.build_shard_dataset(torch.from_numpy(x).cuda(), y, None, chunk_size=1000)
    pre = fitting.RandNysPreconditioner(kern, ds, 64, False, 123, "srht")
    out = {}
    for name, dev_cg in (("device", True), ("host", False)):
        op = fitting.ConjugateGrad()
        op.device_cg = dev_cg
        out[name] = fitting.cg_fit_lib_internal(kern, ds, 1e-8, 500, pre, False, cg_operator=op) + (op.last_fit_info,)
    (w1, n1, l1, info1), (w0, n0, l0, info0) = out["device"], out["host"]
    assert info1["device_cg"] and not info0["device_cg"]
    assert info1["cuda_graph"], "the CG iteration should have been captured in a CUDA graph"
    assert info1["resident"] == (m == 4096)
    # the two loops round differently (dot-product order, the algebraic form of the preconditioner), which shows once the
    # residual nears its float-feature floor: the iteration that crosses the tolerance may differ by one
    # (CG amplifies fp64 rounding differences -- here the atomics order of the matvec and the algebraic form of the
    # preconditioner -- as it loses orthogonality, so only the early history and the solution are comparable)
    assert abs(n1 - n0) <= 2 and len(l1) == n1 and len(l0) == n0
    np.testing.assert_allclose(l1[:8], l0[:8], rtol=1e-6)
    assert l1[-1] < 1e-8 and l0[-1] < 1e-8
    # both solves stop at a 1e-8 relative residual, possibly one iteration apart: the solutions agree to that level
    assert float((w1 - w0).abs().max() / w0.abs().max()) < 1e-6
    # without a preconditioner, and a tolerance it cannot reach in 3 iterations: stops at maxiter, not converged
    op = fitting.ConjugateGrad()
    resid = torch.zeros((m, 2, 1), dtype=torch.float64, device="cuda")
    resid[:, 0, 0] = pre.get_zty() / ds.get_ndatapoints()
    xk, conv, nit, losses = op.fit(ds, kern, None, resid, maxiter=3, tol=1e-30, verbose=False)
    assert not conv and nit == 3 and len(losses) == 3 and np.isfinite(losses).all()


def test_fit_cg_overlapped_build_matches_two_step_fit():
    """fitting.fit_cg (the solve's resident features are built on a side stream under the preconditioner's serial
    factorisation) against the two-step build_preconditioner + CG fit: same iteration count, same weights."""
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(4)
    x = rng.standard_normal((6000, 32)).astype(np.float32)
    y = np.sin(x[:, :4].sum(1)) + 0.1 * rng.standard_normal(6000)
    kern = RBF(x.shape, 4096, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(np.asarray([0.5, 0.3])), logspace=True)
    ds = fitting.build_shard_dataset(torch.from_numpy(x).cuda(), y, None, chunk_size=2048)
    w1, n1, _, pre1 = fitting.fit_cg(kern, ds, 128, "srht", 1e-8, 300)
    pre0 = fitting.RandNysPreconditioner(kern, ds, 128, False, 123, "srht")
    w0, n0, _ = fitting.cg_fit_lib_internal(kern, ds, 1e-8, 300, pre0, False)
    np.testing.assert_allclose(pre1.achieved_ratio, pre0.achieved_ratio, rtol=1e-9)
    assert abs(n1 - n0) <= 1
    assert float((w1 - w0).abs().max() / w0.abs().max()) < 1e-6
    w2, n2, _, _ = fitting.fit_cg(kern, ds, 128, "srht_2", 1e-8, 300)
    assert float((w2 - w0).abs().max() / w0.abs().max()) < 1e-6


def test_fused_predict_matches_materialised():
    """predict() through the fused featurise-dot kernel (Z never written) against (Z * w).sum(1) on transform_x."""
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(8)
    x = rng.standard_normal((5000, 64)).astype(np.float32)
    kern = RBF(x.shape, 2048, random_seed=123, device="cuda")
    kern.set_hyperparams(np.log(np.asarray([0.3, 0.1])), logspace=True)
    w = torch.from_numpy(rng.standard_normal(2048)).cuda()
    pred = fitting.predict(kern, w, x, None, 700, 1.5, 2.0)
    ref = (kern.transform_x(x) @ w).cpu().numpy() * 2.0 + 1.5
    np.testing.assert_allclose(pred, ref, rtol=1e-10, atol=1e-10)


def test_resident_shard_gives_same_weights():
    fitting, g, kern, ds, _, _ = _setup("rbf_f32")
    w0, _, _ = fitting.calc_weights_exact(ds, kern)
    _, _, kern2, ds2, _, _ = _setup("rbf_f32", resident=True)
    w1, _, _ = fitting.calc_weights_exact(ds2, kern2)
    # same features bit for bit; the split-K consumers add partial sums with atomics, so order may differ
    np.testing.assert_allclose(w0.cpu().numpy(), w1.cpu().numpy(), rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("shape", [(300, 64, 1), (1000, 200, 7), (257, 130, 26), (64, 2048, 3)])
def test_consumers_vs_numpy(shape):
    """zty / ztz / atb / cg_matvec on the same Z as NumPy float64: 1e-6 is the north_star bound, 1e-12 is what holds."""
    from xgpr_b200 import ext
    n, m, r = shape
    rng = np.random.default_rng(5)
    z, y, v, a = rng.standard_normal((n, m)), rng.standard_normal(n), rng.standard_normal((m, r)), rng.standard_normal((n, 2 * r + 1))
    Z, Y, V, A = (torch.from_numpy(t).cuda() for t in (z, y, v, a))
    zty = torch.ones(m, dtype=torch.float64, device="cuda")
    ztz = torch.ones((m, m), dtype=torch.float64, device="cuda")
    acc = torch.ones((2 * r + 1, m), dtype=torch.float64, device="cuda")
    w = torch.ones((m, r), dtype=torch.float64, device="cuda")
    ext.zty_accum(Z, Y, zty)
    ext.ztz_accum(Z, ztz)
    ext.atb_accum(A, Z, acc)
    ext.cg_matvec_accum(Z, V, w)
    t = dict(rtol=1e-12, atol=1e-11)
    np.testing.assert_allclose(zty.cpu().numpy(), 1 + z.T @ y, **t)
    np.testing.assert_allclose(ztz.cpu().numpy(), 1 + z.T @ z, **t)
    np.testing.assert_allclose(acc.cpu().numpy(), 1 + a.T @ z, **t)
    np.testing.assert_allclose(w.cpu().numpy(), 1 + z.T @ (z @ v), **t)


@pytest.mark.parametrize("dp", [False, True])
@pytest.mark.parametrize("case", [(700, 64, 2048, True), (333, 20, 100, True), (129, 7, 6, False), (260, 512, 2048, True),
                                  (50, 130, 1000, False), (1000, 64, 16384, True), (40, 1000, 300, True),
                                  (5000, 64, 16384, True), (777, 128, 16384, False), (300, 128, 32768, True),
                                  (64, 9, 64, True), (100, 300, 8192, True)])
def test_fused_matvec_equals_materialised(case, dp):
    """rbf_zv / rbf_ztv (features regenerated on chip) against transform_x(x) @ v and transform_x(x).T @ t."""
    from xgpr_b200.kernels import RBF, Matern
    n, d, m, icpt = case
    rng = np.random.default_rng(11)
    cls = Matern if d == 20 else RBF
    parms = {"intercept": icpt}
    if cls is Matern:
        parms["matern_nu"] = 1.5
    k = cls((n, d), m, device="cuda", double_precision=dp, kernel_spec_parms=parms)
    k.set_hyperparams(np.log(np.asarray([0.3, 1.0 / np.sqrt(d)])), logspace=True)
    x = rng.standard_normal((n, d))
    z = k.transform_x(x)
    v = torch.from_numpy(rng.standard_normal(m)).cuda()
    tin = torch.from_numpy(rng.standard_normal(n)).cuda()
    xs = k.stage_scaled(x)
    t = torch.full((n,), 0.5, dtype=torch.float64, device="cuda")
    w = torch.full((m,), -0.25, dtype=torch.float64, device="cuda")
    k.fused_zv(xs, v, t)
    k.fused_ztv(xs, tin, w)
    np.testing.assert_allclose(t.cpu().numpy(), 0.5 + (z @ v).cpu().numpy(), rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(w.cpu().numpy(), -0.25 + (z.T @ tin).cpu().numpy(), rtol=1e-11, atol=1e-12)
    # one-pass w += Z^T (Z v), wherever a row's repeats fit one CTA
    from xgpr_b200 import ext
    assert k.fused_zzv_ok() == ext.rbf_zzv_supported(d, k.num_freqs)
    if k.fused_zzv_ok():
        w2 = torch.full((m,), 0.125, dtype=torch.float64, device="cuda")
        k.fused_zzv(xs, v, w2)
        ref = 0.125 + (z.T @ (z @ v)).cpu().numpy()
        np.testing.assert_allclose(w2.cpu().numpy(), ref, rtol=1e-10, atol=1e-11 * float(np.abs(ref).max()))
    else:
        with pytest.raises(RuntimeError):
            k.fused_zzv(xs, v, torch.zeros(m, dtype=torch.float64, device="cuda"))


@pytest.mark.parametrize("r", [1, 3, 24, 25, 26, 27, 32, 40, 58])
@pytest.mark.parametrize("case", [(700, 64, 2048, True), (333, 20, 100, False), (2100, 128, 32768, True)])
def test_encoded_multi_rhs_matvec_equals_fp64_matvec(case, r):
    """xgpr_b200_cg_matvec_accum_c32 (DMMA kernels reading the float32 encoding of Z, 26-column tile for SLQ's 25 probes +
    y) against xgpr_b200_cg_matvec_accum_f64 on transform_x's Z and against NumPy.  Z = double(C) * scale exactly, so the
    only differences are where the scale is applied and the summation order: rtol 1e-11 of the result's scale."""
    from xgpr_b200 import ext
    from xgpr_b200.kernels import RBF
    n, d, m, icpt = case
    rng = np.random.default_rng(11)
    x = rng.standard_normal((n, d)).astype(np.float32)
    kern = RBF(x.shape, m, random_seed=123, device="cuda")
    kern.fit_intercept = icpt
    kern.set_hyperparams(np.log(np.asarray([1.0, 0.3])), logspace=True)
    assert kern.encoding_ok()
    z = kern.transform_x(x)
    c = torch.empty((n, m), dtype=torch.float32, device="cuda")
    kern.resident_generate(x, c)
    if icpt:
        c[:, 0] = 1.0
    scale = kern.feature_scale()
    v = torch.from_numpy(rng.standard_normal((m, r))).cuda()
    vs = v * scale
    if icpt:
        vs[0] = v[0]
    w0 = torch.ones((m, r), dtype=torch.float64, device="cuda")
    w1 = torch.ones((m, r), dtype=torch.float64, device="cuda")
    ext.cg_matvec_accum(z, v, w0)
    ext.cg_matvec_accum_c32(c, vs, w1, scale, icpt)
    zh = z.cpu().numpy()
    want = 1.0 + zh.T @ (zh @ v.cpu().numpy())
    tol = 1e-11 * float(np.abs(want).max())
    np.testing.assert_allclose(w0.cpu().numpy(), want, rtol=1e-10, atol=tol)
    np.testing.assert_allclose(w1.cpu().numpy(), want, rtol=1e-10, atol=tol)


def test_multi_rhs_tail_tile_equals_full_tile(monkeypatch):
    """25 / 26 right-hand sides: the 24 + 2 tile (columns 24, 25 on DFMA) against NumPy, fp64 and encoded inputs, with a row
    count and a feature count that are not multiples of the tile."""
    from xgpr_b200 import ext
    rng = np.random.default_rng(12)
    for n, m, r in ((1000, 520, 26), (130, 4100, 25), (4096, 1024, 26)):
        z = rng.standard_normal((n, m))
        v = rng.standard_normal((m, r))
        want = z.T @ (z @ v)
        w = torch.zeros((m, r), dtype=torch.float64, device="cuda")
        ext.cg_matvec_accum(torch.from_numpy(z).cuda(), torch.from_numpy(v).cuda(), w)
        np.testing.assert_allclose(w.cpu().numpy(), want, rtol=1e-11, atol=1e-11 * float(np.abs(want).max()))
        if m % 4 == 0:
            zf = z.astype(np.float32)
            wantf = 0.25 * (zf.astype(np.float64).T @ (zf.astype(np.float64) @ v))
            w = torch.zeros((m, r), dtype=torch.float64, device="cuda")
            ext.cg_matvec_accum_c32(torch.from_numpy(zf).cuda(), torch.from_numpy(v * 0.5).cuda(), w, 0.5, False)
            np.testing.assert_allclose(w.cpu().numpy(), wantf, rtol=1e-11, atol=1e-11 * float(np.abs(wantf).max()))


@pytest.mark.parametrize("kname", ["RBF", "MiniARD"])
def test_multi_rhs_cg_encoded_chunks_match_fp64_chunks(kname, monkeypatch):
    """A 26-RHS solve (the shape of approximate_nmll's) through (a) fp64 feature chunks, (b) the float32 encoding with every
    chunk kept, (c) the encoding with room for only SOME chunks (the others regenerated per iteration), (d) nothing kept."""
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import KERNEL_NAME_TO_CLASS
    rng = np.random.default_rng(3)
    n, d, m = 3000, 20, 512
    x = rng.standard_normal((n, d)).astype(np.float32)
    y = np.sin(x[:, :3].sum(1)) + 0.1 * rng.standard_normal(n)
    parms = {"split_points": [7]} if kname == "MiniARD" else {}
    kern = KERNEL_NAME_TO_CLASS[kname](x.shape, m, random_seed=123, device="cuda", kernel_spec_parms=parms)
    hp = [2.0, 0.4, 0.25] if kname == "MiniARD" else [2.0, 0.4]        # lambda = 2: unpreconditioned CG converges
    kern.set_hyperparams(np.log(np.asarray(hp)), logspace=True)
    ds = fitting.build_shard_dataset(x, y, None, chunk_size=250)
    monkeypatch.setattr(fitting, "_work_rows", lambda kernel, budget_bytes=0, cap=0: 512)     # 6 work chunks
    rhs = torch.from_numpy(rng.standard_normal((m, 26))).cuda()
    outs = []
    for mode in ("f64", "all", "some", "none"):
        cg = fitting.ConjugateGrad()
        cg.encoded_multi_rhs = mode != "f64"
        if mode == "some":                       # room for the work buffer + 3 of the 6 chunks
            cg.encoded_fraction = (4.5 * 512 * m * 4) / fitting._available_hbm(torch.device("cuda", torch.cuda.current_device()))
        cg.keep_resident = mode != "none"
        resid = torch.zeros((m, 2, 26), dtype=torch.float64, device="cuda")
        resid[:, 0, :] = rhs
        xk, conv, nit, _ = cg.fit(ds, kern, None, resid, maxiter=300, tol=1e-11, verbose=False)
        assert conv, (mode, nit)
        if mode != "f64":
            kept = len(cg._encoded[1]["kept"])
            assert kept == {"all": 6, "some": 3, "none": 0}[mode], (mode, kept)
        outs.append(xk.cpu().numpy())
    for o in outs[1:]:
        # converged solutions of the same system: CG amplifies the rounding differences of the two forms by the condition
        # number, so the bound is on the solution's scale (the north_star weight tolerance is 1e-6)
        np.testing.assert_allclose(o, outs[0], rtol=0, atol=1e-8 * float(np.abs(outs[0]).max()))


def _guarded(shape, fill=0.0, pad=2048):
    """A float64 CUDA view of `shape` with NaN guard bands on both sides of its storage."""
    n = int(np.prod(shape))
    buf = torch.full((n + 2 * pad,), float("nan"), dtype=torch.float64, device="cuda")
    view = buf[pad:pad + n].view(shape)
    view.fill_(fill)
    return buf, view, pad, n


@pytest.mark.parametrize("shape", [(130, 4100, 25), (1000, 520, 26), (257, 132, 3), (33, 8, 1), (4097, 260, 32), (64, 16388, 26)])
def test_consumer_outputs_stay_inside_their_arrays(shape):
    """Guard bands around every array the multi-RHS / single-RHS consumers WRITE (w, the (n, r) intermediate, Z^T y): no
    element outside the arrays may change, for row / feature / RHS counts that are not multiples of any tile.  (The pool's
    compute-sanitizer is closed, so out-of-bounds stores are looked for this way.)"""
    from xgpr_b200 import ext
    n, m, r = shape
    rng = np.random.default_rng(21)
    zf = rng.standard_normal((n, m)).astype(np.float32)
    z = torch.from_numpy(zf.astype(np.float64)).cuda()
    c = torch.from_numpy(zf).cuda()
    v = torch.from_numpy(rng.standard_normal((m, r))).cuda()
    y = torch.from_numpy(rng.standard_normal(n)).cuda()
    want = zf.astype(np.float64).T @ (zf.astype(np.float64) @ v.cpu().numpy())
    tol = dict(rtol=1e-10, atol=1e-11 * float(np.abs(want).max()))
    for which in ("f64", "c32"):
        wb, w, pad, cnt = _guarded((m, r))
        sb, s, spad, scnt = _guarded((n, r), fill=7.0)
        if which == "f64":
            ext.cg_matvec_accum(z, v, w, s)
        else:
            ext.cg_matvec_accum_c32(c, v, w, 1.0, False, s)
        torch.cuda.synchronize()
        np.testing.assert_allclose(w.cpu().numpy(), want, **tol)
        for buf, p, k in ((wb, pad, cnt), (sb, spad, scnt)):
            assert bool(torch.isnan(buf[:p]).all()) and bool(torch.isnan(buf[p + k:]).all()), (which, shape)
    tb, t, pad, cnt = _guarded((m,))
    ext.zty_accum(z, y, t)
    torch.cuda.synchronize()
    wanty = zf.astype(np.float64).T @ y.cpu().numpy()
    np.testing.assert_allclose(t.cpu().numpy(), wanty, rtol=1e-10, atol=1e-11 * float(np.abs(wanty).max()))
    assert bool(torch.isnan(tb[:pad]).all()) and bool(torch.isnan(tb[pad + cnt:]).all())


@pytest.mark.parametrize("icpt", [True, False])
@pytest.mark.parametrize("shape", [(700, 64, 2048, 64), (2100, 20, 4096, 512), (333, 128, 16384, 512), (150, 8, 1024, 100)])
def test_encoded_preconditioner_steps_equal_fp64_steps(shape, icpt):
    """The preconditioner's accumulation pass fed from the float32 encoding of Z (xgpr_b200_zty_accum_c32,
    xgpr_b200_srht_sample_c32, xgpr_b200_atb_accum_tc_c32) against the same steps on fp64 Z = transform_x(x): the SRHT sample
    is bit-identical (same values into the same butterflies), Z^T y and the sketch product agree to rounding."""
    from xgpr_b200 import ext
    from xgpr_b200.kernels import RBF, SRHTCompressor
    n, d, m, rank = shape
    rng = np.random.default_rng(31)
    x = rng.standard_normal((n, d)).astype(np.float32)
    y = torch.from_numpy(rng.standard_normal(n)).cuda()
    kern = RBF(x.shape, m, random_seed=123, device="cuda")
    kern.fit_intercept = icpt
    kern.set_hyperparams(np.log(np.asarray([1.0, 0.3])), logspace=True)
    z = kern.transform_x(x)
    c = torch.empty((n, m), dtype=torch.float32, device="cuda")
    kern.resident_generate(x, c)
    if icpt:
        c[:, 0] = 1.0
    scale = kern.feature_scale()
    comp = SRHTCompressor(rank, m, random_seed=123, device="cuda")
    s64 = comp.transform_x(z)
    s32 = comp.transform_encoded(c, scale, icpt)
    assert s32 is not None and torch.equal(s64, s32)
    t64 = torch.zeros(m, dtype=torch.float64, device="cuda")
    t32 = torch.zeros(m, dtype=torch.float64, device="cuda")
    ext.zty_accum(z, y, t64)
    ext.zty_accum_c32(c, y, t32, scale, icpt)
    np.testing.assert_allclose(t32.cpu().numpy(), t64.cpu().numpy(), rtol=1e-12, atol=1e-13 * float(t64.abs().max()))
    a64 = torch.zeros((rank, m), dtype=torch.float64, device="cuda")
    a32 = torch.zeros((rank, m), dtype=torch.float64, device="cuda")
    ext.atb_accum_tc(s64, z, a64)
    ext.atb_accum_tc_c32(s64, c, a32, scale, icpt)
    want = s64.cpu().numpy().T @ z.cpu().numpy()
    tol = 1e-10 * float(np.abs(want).max())
    np.testing.assert_allclose(a64.cpu().numpy(), want, rtol=0, atol=tol)
    np.testing.assert_allclose(a32.cpu().numpy(), want, rtol=0, atol=tol)


def test_fit_cg_with_encoded_preconditioner_pass_matches_fp64_pass(monkeypatch):
    """fitting.fit_cg end to end with the preconditioner's accumulation pass reading the resident float32 encoding (default
    when the solve keeps it) against XGPR_B200_PRECOND_F64=1: same ratio, iteration count and weights."""
    from xgpr_b200 import fitting
    from xgpr_b200.kernels import RBF
    rng = np.random.default_rng(41)
    n, d, m = 6000, 20, 4096
    x = rng.standard_normal((n, d)).astype(np.float32)
    y = np.sin(x[:, :3].sum(1)) + 0.1 * rng.standard_normal(n)
    outs = []
    for f64 in ("1", "0"):
        monkeypatch.setenv("XGPR_B200_PRECOND_F64", f64)
        kern = RBF(x.shape, m, random_seed=123, device="cuda")
        kern.set_hyperparams(np.log(np.asarray([0.3, 0.4])), logspace=True)
        ds = fitting.build_shard_dataset(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), None, chunk_size=2000)
        w, nit, _, pre = fitting.fit_cg(kern, ds, 256, "srht", 1e-8, 300, 123, False)
        outs.append((w.cpu().numpy(), nit, float(pre.achieved_ratio), pre.get_zty().cpu().numpy()))
    (w0, n0, r0, t0), (w1, n1, r1, t1) = outs
    assert abs(n0 - n1) <= 1
    np.testing.assert_allclose(r1, r0, rtol=1e-7)
    np.testing.assert_allclose(t1, t0, rtol=1e-12, atol=1e-13 * float(np.abs(t0).max()))
    np.testing.assert_allclose(w1, w0, rtol=0, atol=1e-7 * float(np.abs(w0).max()))
