"""The NMLL call sites of the hot path: the callers of the Gram / gradient-feature / multi-RHS CG
consumers, mirrored on torch tensors and the C-ABI kernels.

  exact_nmll                 src/xGPR/xgp_regression.py:150-205
  calc_gradient_terms        src/xGPR/scoring_toolkit/nmll_gradient_tools.py:12-90
  exact_nmll_reg_grad        src/xGPR/scoring_toolkit/nmll_gradient_tools.py:94-161 (+ gpu_cho_calcs :179-206)
  exact_nmll_gradient        src/xGPR/xgp_regression.py:209-260
  approximate_nmll           src/xGPR/xgp_regression.py:264-377   (stochastic Lanczos quadrature, 25 probes + Z^T y)
  generate_normal_probes     src/xGPR/scoring_toolkit/probe_generators.py:9-31
  estimate_logdet            src/xGPR/scoring_toolkit/approximate_nmll_calcs.py:12-50
  optimize_alpha_beta        src/xGPR/scoring_toolkit/alpha_beta_optimizer.py:13-39

Rows shard across GPUs as in fitting.py: every accumulation (Z^T Z, Z^T y, dZ^T y, dZ^T Z, the 26-RHS CG matvec) is a
per-rank partial sum followed by ONE fp64 all-reduce; the M x M factorisations are replicated.  The Gram-shaped
accumulations (Z^T Z and dZ_i^T Z) run on the tensor cores (csrc/gram_tc_kernels.cu) unless gram="f64".
"""
import warnings

import numpy as np
import torch
from scipy.linalg import eigh_tridiagonal

from . import ext as _cuda_ops
from . import fitting
from .fitting import RowShardComm, _device_of, _new

MAX_RATIO_CHECK_RFFS = 8192                          # model_baseclass.py:470
DEFAULT_SCORE_IF_PROBLEM = 1e40                      # constants.py:12
DEFAULT_NMLL_PARAMS = {"max_rank": 1024, "preconditioner_mode": "srht_2", "nsamples": 25, "nmll_iter": 500,
                       "nmll_tol": 1e-6}             # constants.py:15-16


def optimize_alpha_beta(lambda_, nll_terms, ndatapoints, nrffs, beta_max=10., beta_min=0.1):
    """lambda = alpha / beta; picks the beta in [beta_min, beta_max] that minimises the NMLL for this lambda and returns
    (score, beta).  alpha_beta_optimizer.py:13-39."""
    beta = np.sqrt(2 * nll_terms[0] / (ndatapoints * lambda_ ** 2))
    beta = max(min(beta, beta_max), beta_min)
    score = nll_terms[0] / (beta * lambda_) ** 2 + (ndatapoints - nrffs) * np.log(lambda_)
    score += nll_terms[1] + ndatapoints * np.log(beta)
    return score + 0.5 * ndatapoints * np.log(2 * np.pi), beta


# ----------------------------------------------------------------------------------------- exact NMLL
def exact_nmll(hyperparams, dataset, kernel, comm=None, ops=_cuda_ops, gram="tc", gram_slices=(5, 8)):
    """xgp_regression.py:150-205: NMLL from the Cholesky factor of Z^T Z + lambda^2 I."""
    kernel.set_hyperparams(hyperparams, logspace=True)
    ndatapoints = dataset.get_ndatapoints()
    z_trans_z, z_trans_y, y_trans_y = fitting.calc_design_mat(dataset, kernel, comm, ops, gram, gram_slices)
    try:
        chol, weights = fitting.direct_weight_calc(z_trans_z, z_trans_y, kernel)
    except Exception:
        warnings.warn(f"Near-singular matrix encountered when calculating score for hyperparameters {hyperparams}.")
        return DEFAULT_SCORE_IF_PROBLEM
    nll1 = float(0.5 * (y_trans_y - z_trans_y @ weights))
    nll2 = float(torch.log(chol.diagonal()).sum())
    negloglik, _ = optimize_alpha_beta(kernel.get_lambda(), np.array([nll1, nll2]), ndatapoints, kernel.get_num_rffs())
    if np.isnan(negloglik):
        warnings.warn(f"Near-singular matrix encountered when calculating score for hyperparameters {hyperparams}.")
        return DEFAULT_SCORE_IF_PROBLEM
    return float(negloglik)


def calc_gradient_terms(dataset, kernel, subsample=1, comm=None, ops=_cuda_ops, gram="tc", gram_slices=(5, 8)):
    """nmll_gradient_tools.py:12-90.  Per chunk: Z, dZ/dsigma_i, then Z^T Z, Z^T y, y^T y, dZ_i^T y and dZ_i^T Z
    (one more M x M contraction per kernel-specific hyperparameter); inner_deriv is symmetrised at the end."""
    if subsample > 1 or subsample < 0.01:
        raise RuntimeError("Subsample must be in the range [0.01, 1].")
    comm = comm or RowShardComm()
    dev = _device_of(kernel)
    m = kernel.get_num_rffs()
    nh = kernel.get_hyperparams().shape[0] - 1
    z_trans_z, z_trans_y, y_trans_y = _new((m, m), dev), _new((m,), dev), _new((1,), dev)
    dz_dsigma_ty = _new((m, nh), dev)
    inner = [_new((m, m), dev) for _ in range(nh)]
    accumulate, finish = fitting._gram_accumulator(ops, gram, gram_slices)
    atb = (getattr(ops, "atb_accum_tc", None) if gram == "tc" else None) or ops.atb_accum
    nlocal = 0
    rng = np.random.default_rng(123)
    for xin, yin, ldata in dataset.get_chunked_data():
        if subsample != 1:
            idx_size = max(1, int(subsample * xin.shape[0]))
            idx = rng.choice(xin.shape[0], idx_size, replace=False)
            tidx = torch.as_tensor(idx, device=xin.device) if isinstance(xin, torch.Tensor) else idx
            xin = xin[tidx, ...]
            yin = yin[torch.as_tensor(idx, device=yin.device)] if isinstance(yin, torch.Tensor) else yin[idx]
            ldata = None if ldata is None else ldata[idx]
        z, dz, y = kernel.gradient_x_y(xin, yin, ldata)
        y = y.to(torch.float64)
        ops.zty_accum(z, y, z_trans_y)
        accumulate(z, z_trans_z)
        y_trans_y += y @ y
        nlocal += z.shape[0]
        for i in range(dz.shape[2]):
            dzi = dz[:, :, i].contiguous()
            col = _new((m,), dev)
            ops.zty_accum(dzi, y, col)
            dz_dsigma_ty[:, i] += col
            atb(dzi, z, inner[i])
    finish(z_trans_z)
    comm.all_reduce_(z_trans_z)
    nd = _new((1,), dev)
    nd += nlocal
    comm.all_reduce_(z_trans_y, y_trans_y, dz_dsigma_ty, nd)
    for mat in inner:
        comm.all_reduce_(mat)
    inner_deriv = torch.stack([mat + mat.T for mat in inner], dim=2) if nh > 0 else _new((m, m, 0), dev)
    return z_trans_z, z_trans_y, float(y_trans_y[0]), dz_dsigma_ty, inner_deriv, int(round(float(nd[0])))


def _cho_solve(chol, target):
    """(L L^T)^-1 target by two triangular solves (gpu_cho_solver, nmll_gradient_tools.py:170-177)."""
    two_d = target.dim() == 2
    t = target if two_d else target[:, None]
    sol = torch.linalg.solve_triangular(chol, t, upper=False)
    sol = torch.linalg.solve_triangular(chol.T, sol, upper=True)
    return sol if two_d else sol[:, 0]


def exact_nmll_reg_grad(z_trans_z, z_trans_y, y_trans_y, hparams, ndatapoints, dz_dsigma_ty, inner_deriv):
    """NMLL and its gradient w/r/t the LOG hyperparameters.  nmll_gradient_tools.py:94-161; z_trans_z is modified in
    place exactly as there (lambda^2 added, then removed again after the factorisation).

    Same quantities, different algebra for the two O(M^3) pieces the reference flags "TODO: Optimize" (:172, :181-183).
    With A = Z^T Z + lambda^2 I = L L^T the reference forms L^-1 explicitly for ||L^-1||_F^2 (:205, :151) and, per
    kernel hyperparameter, solves A X = D_i with M right-hand sides only to take trace(X) (:155, :159) -- 1/3 + 2 nh
    M^3-sized triangular solves.  Both are traces against A^-1:
        ||L^-1||_F^2 = trace(A^-1),        trace(A^-1 D_i) = sum(A^-1 o D_i)   (D_i = dZ_i^T Z + Z^T dZ_i is symmetric),
    so ONE Cholesky inverse (potri, 2/3 M^3) serves every hyperparameter and the per-hyperparameter work drops to an
    O(M^2) elementwise sum: 2.7x fewer flops for one lengthscale, 4.7x for MiniARD with two."""
    m = z_trans_z.shape[0]
    lam = float(hparams[0])
    z_trans_z.diagonal().add_(lam ** 2)
    chol = torch.linalg.cholesky(z_trans_z)
    weights = _cho_solve(chol, z_trans_y)
    z_trans_z.diagonal().sub_(lam ** 2)
    a_inv = torch.cholesky_inverse(chol)
    nll1 = float(0.5 * (y_trans_y - z_trans_y @ weights))
    nll2 = float(torch.log(chol.diagonal()).sum())
    negloglik, beta = optimize_alpha_beta(lam, np.array([nll1, nll2]), float(ndatapoints), float(m))

    grad = np.zeros((hparams.shape[0]))
    alpha = lam * beta
    dnll_dlambda = (1 / (beta ** 2 * lam ** 3)) * (float(z_trans_y @ weights) - y_trans_y)
    dnll_dlambda += (1 / (beta ** 2 * lam)) * float(weights @ weights)
    dnll_dlambda += (ndatapoints - m) / lam
    dnll_dlambda += lam * float(a_inv.diagonal().sum())
    grad[0] = float(dnll_dlambda)
    for i in range(grad.shape[0] - 1):
        d_i = inner_deriv[:, :, i]
        dnll_dsigma = -2 * float(weights @ dz_dsigma_ty[:, i])
        dnll_dsigma += float(weights @ (d_i @ weights))
        dnll_dsigma *= (0.5 / alpha ** 2)
        dnll_dsigma += 0.5 * float((a_inv * d_i).sum())
        grad[i + 1] = float(dnll_dsigma)
    grad *= hparams
    return negloglik, grad, beta


def exact_nmll_gradient(hyperparams, dataset, kernel, subsample=1, comm=None, ops=_cuda_ops, gram="tc", gram_slices=(5, 8)):
    """xgp_regression.py:209-260."""
    init_hparams = kernel.get_hyperparams()
    kernel.set_hyperparams(hyperparams, logspace=True)
    hparams = kernel.get_hyperparams(logspace=False)
    z_trans_z, z_trans_y, y_trans_y, dz_dsigma_ty, inner_deriv, nsamples = calc_gradient_terms(
        dataset, kernel, subsample, comm, ops, gram, gram_slices)
    try:
        negloglik, grad, _ = exact_nmll_reg_grad(z_trans_z, z_trans_y, y_trans_y, hparams, nsamples, dz_dsigma_ty,
                                                 inner_deriv)
    except Exception:
        return DEFAULT_SCORE_IF_PROBLEM, hyperparams - init_hparams
    if np.isnan(negloglik):
        return DEFAULT_SCORE_IF_PROBLEM, hyperparams - init_hparams
    return float(negloglik), grad


# ----------------------------------------------------------------------------------------- approximate NMLL (SLQ)
def generate_normal_probes(nsamples, num_rffs, random_seed=123, preconditioner=None, device=None):
    """Host-drawn standard-normal probes, mapped through P^(1/2) when preconditioned.  probe_generators.py:9-31."""
    rng = np.random.default_rng(random_seed)
    probes = torch.from_numpy(rng.standard_normal(size=(num_rffs, nsamples)))
    if device is not None:
        probes = probes.to(device)
    if preconditioner is not None:
        probes = preconditioner.matvec_for_sampling(probes)
    return probes


def estimate_logdet(alphas, betas, num_rffs, preconditioner=None):
    """log det (Z^T Z + lambda^2 I) from the CG coefficients of the probe systems: per probe the Lanczos tridiagonal
    T (diag 1/alpha_k + beta_{k-1}/alpha_{k-1}, off-diag sqrt(beta_k)/alpha_k), Gauss quadrature of log with the squared
    first components of T's eigenvectors as weights.  approximate_nmll_calcs.py:12-50 (host, scipy stev)."""
    mat_diag = 1 / alphas
    mat_diag[1:, :] += betas[:-1, :] / alphas[:-1, :]
    upper_diag = np.sqrt(betas) / alphas
    logdets = np.zeros((mat_diag.shape[1]))
    for i in range(mat_diag.shape[1]):
        eigvals, eigvecs = eigh_tridiagonal(mat_diag[:, i], upper_diag[:-1, i], lapack_driver="stev")
        weights = eigvecs[0, :] ** 2
        logdets[i] += (weights * np.log(eigvals)).sum()
    logdet = num_rffs * logdets.sum() / alphas.shape[1]
    if preconditioner is not None:
        logdet += preconditioner.get_logdet()
    return float(logdet)


def check_rank_ratio(dataset, kernel, sample_frac, max_rank, random_seed=123, comm=None, ops=_cuda_ops,
                     compressor_cls=None):
    """min_eig / lambda^2 / sample_frac of a rank-`max_rank` Nystrom approximation built from a random `sample_frac` of
    every chunk (a per-chunk permutation prefix).  model_baseclass.py:438-475, rand_nys_constructors.py:60-93, :301-360."""
    if sample_frac < 0.01 or sample_frac > 1:
        raise RuntimeError("sample_frac must be in the range [0.01, 1]")
    comm = comm or RowShardComm()
    if compressor_cls is None:
        from .kernels import SRHTCompressor as compressor_cls
    dev = _device_of(kernel)
    # the reference runs this pass on min(num_rffs, 8192) random features: `self.num_rffs = min(self.num_rffs, 8192)`
    # re-initialises the kernel (same class, seed, settings, hyperparameters) before srht_ratio_check and restores it
    # afterwards (model_baseclass.py:465-474).  Same here, on a temporary kernel; the caller's kernel is untouched.
    if kernel.get_num_rffs() > MAX_RATIO_CHECK_RFFS and hasattr(kernel, "with_num_rffs"):
        kernel = kernel.with_num_rffs(MAX_RATIO_CHECK_RFFS)
    m = kernel.get_num_rffs()
    acc = _new((max_rank, m), dev)
    compressor = compressor_cls(max_rank, m, random_seed=random_seed, device=kernel.device)
    atb = getattr(ops, "atb_accum_tc", None) or ops.atb_accum
    rng = np.random.default_rng(random_seed)
    for xdata, ldata in dataset.get_chunked_x_data():
        cutoff = max(int(sample_frac * float(xdata.shape[0])), 1)
        idx = rng.permutation(xdata.shape[0])[:cutoff]
        xs = xdata[torch.as_tensor(idx, device=xdata.device), ...] if isinstance(xdata, torch.Tensor) else xdata[idx, ...]
        z = kernel.transform_x(xs, None if ldata is None else ldata[idx])
        atb(compressor.transform_x(z), z, acc)
    comm.all_reduce_(acc)
    c_s1, c_v1 = fitting._psd_eig(compressor.transform_x(acc))
    mask = c_s1 < 1e-14
    c_s1 = 1.0 / torch.sqrt(c_s1.clamp(min=1e-14))
    c_s1[mask] = 0
    _, s = fitting._tall_svd(acc.T @ c_v1.T @ (c_s1[:, None] * c_v1))
    return float((s ** 2).min() / kernel.get_lambda() ** 2) / sample_frac


def autoselect_preconditioner(dataset, kernel, min_rank=512, max_rank=3000, increment_size=512, always_use_srht2=False,
                              ratio_target=30., random_seed=123, comm=None, ops=_cuda_ops, compressor_cls=None):
    """model_baseclass.py:376-434: grow the rank in steps until the estimated ratio min_eig / lambda^2 drops under the
    target, falling back to "srht_2" at max_rank."""
    sample_frac, method, ratio, rank = 0.2, "srht", np.inf, min_rank
    m = kernel.get_num_rffs()
    if rank >= m:
        rank, ratio = m - 1, 0.5 * ratio_target
    if dataset.get_ndatapoints() < 5000:
        sample_frac = 1
    while ratio > ratio_target and rank < max_rank:
        ratio = check_rank_ratio(dataset, kernel, sample_frac, rank, random_seed, comm, ops, compressor_cls)
        if ratio > ratio_target:
            if (rank + increment_size) < max_rank and (rank + increment_size) < m:
                rank += increment_size
            else:
                rank = max_rank
                if rank > m:
                    rank = m - 1
                method = "srht_2"
                break
    if always_use_srht2:
        method = "srht_2"
    return fitting.RandNysPreconditioner(kernel, dataset, rank, False, random_seed, method, comm, ops, compressor_cls)


# phases of the last approximate_nmll call on this process (device time, CUDA events): for bench.py / tools
LAST_NMLL_INFO = {}


def approximate_nmll(hyperparams, dataset, kernel, manual_settings=None, random_seed=123, comm=None, ops=_cuda_ops,
                     compressor_cls=None):
    """xgp_regression.py:264-377: preconditioned CG on nsamples + 1 right-hand sides (Z^T y / N and the probes), log det by
    stochastic Lanczos quadrature from the CG coefficients, data-fit term from the first solution column."""
    comm = comm or RowShardComm()
    kernel.set_hyperparams(hyperparams, logspace=True)
    settings = dict(DEFAULT_NMLL_PARAMS)
    m = kernel.get_num_rffs()
    timed = kernel.device == "cuda" and torch.cuda.is_available()
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if timed else None
    if timed:
        marks[0].record()
    if manual_settings is not None:
        for key in settings:
            if key in manual_settings:
                settings[key] = manual_settings[key]
        if settings["max_rank"] >= m:
            settings["max_rank"] = m - 1
        preconditioner = fitting.RandNysPreconditioner(kernel, dataset, settings["max_rank"], False, random_seed,
                                                       settings["preconditioner_mode"], comm, ops, compressor_cls)
    else:
        preconditioner = autoselect_preconditioner(dataset, kernel, 512, 3000, always_use_srht2=True,
                                                   random_seed=random_seed, comm=comm, ops=ops,
                                                   compressor_cls=compressor_cls)
    z_trans_y, y_trans_y = preconditioner.get_zty(), preconditioner.get_yty()
    dev = z_trans_y.device
    if timed:
        marks[1].record()
    resid = torch.zeros((m, 2, settings["nsamples"] + 1), dtype=torch.float64, device=dev)
    resid[:, 0, 0] = z_trans_y / dataset.get_ndatapoints()
    resid[:, 0, 1:] = generate_normal_probes(settings["nsamples"], m, random_seed, preconditioner, dev)
    cg_operator = fitting.ConjugateGrad(comm, ops)
    x_k, alphas, betas = cg_operator.fit(dataset, kernel, preconditioner, resid, settings["nmll_iter"],
                                         settings["nmll_tol"], verbose=False, nmll_settings=True)
    x_k[:, 0] *= dataset.get_ndatapoints()
    if timed:
        marks[2].record()
        marks[2].synchronize()
        enc = cg_operator._encoded[1] if getattr(cg_operator, "_encoded", None) else None
        LAST_NMLL_INFO.clear()
        LAST_NMLL_INFO.update({"precond_s": marks[0].elapsed_time(marks[1]) * 1e-3,
                               "cg_s": marks[1].elapsed_time(marks[2]) * 1e-3, "cg_iters": int(alphas.shape[0]),
                               "rhs": int(resid.shape[2]),
                               "encoded_rows_kept": None if enc is None else int(sum(c.shape[0] for c in enc["kept"].values())),
                               "rows_local": int(dataset.get_local_ndatapoints())})
    logdet = estimate_logdet(alphas, betas, m, preconditioner)
    nll1 = float(0.5 * (y_trans_y - z_trans_y @ x_k[:, 0]))
    negloglik, _ = optimize_alpha_beta(kernel.get_lambda(), np.array([nll1, 0.5 * logdet]),
                                       dataset.get_ndatapoints(), m)
    return float(negloglik)
