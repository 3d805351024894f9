"""The L4 call sites of the hot path: chunk loop -> kernel.transform_x -> accumulate.

Mirrors, on torch tensors and the C-ABI consumers of Z (xgpr_b200_{zty,ztz,cg_matvec}_accum_f64):

  calc_zty / calc_design_mat / direct_weight_calc / calc_var_design_mat
        src/xGPR/scoring_toolkit/exact_nmll_calcs.py:13-38, :42-78, :82-112, :116-139
  calc_weights_exact / calc_variance_exact
        src/xGPR/fitting_toolkit/exact_fitting_toolkit.py:16-36, :40-68   (incl. lambda^2 added twice)
  ConjugateGrad / cg_fit_lib_internal
        src/xGPR/fitting_toolkit/cg_tools.py:26-156, cg_fitting_toolkit.py:18-70
  RandNysPreconditioner (methods "srht", "srht_2", ...)
        src/xGPR/preconditioners/rand_nys_constructors.py:18-123, :128-296,
        src/xGPR/preconditioners/rand_nys_preconditioners.py:18-72
  predict   src/xGPR/xgp_regression.py:105-134

Rows shard across GPUs: every rank holds a contiguous row block in its own
``ShardDataset``; parameters are replicated; ``RowShardComm.all_reduce_`` (NCCL, fp64 sum) runs at the
reduction points only -- Z^T y / y^T y, Z^T Z, the Nystrom accumulator, and the CG matvec partial sum
once per iteration.  Dense factorizations (Cholesky, SVD, QR) are replicated library calls, as in the
reference.

``ops`` is the provider of the three consumers; the default is the CUDA library (``xgpr_b200.ext``) and
fails loudly on non-CUDA tensors.  Tests inject an oracle-backed provider to run the host logic under
``gloo`` on CPU.
"""
import os
import warnings

import numpy as np
import torch

from . import ext as _cuda_ops


# ----------------------------------------------------------------------------------------- plumbing
class RowShardComm:
    """Thin wrapper over torch.distributed for the row-sharded reductions (no-op when world size is 1)."""

    def __init__(self, group=None, single=False):
        import torch.distributed as dist
        self.dist = dist if (dist.is_available() and dist.is_initialized() and not single) else None
        self.group = group
        self.world = self.dist.get_world_size(group) if self.dist else 1
        self.rank = self.dist.get_rank(group) if self.dist else 0
        self.n_allreduce = 0
        self.bytes_allreduce = 0

    def all_reduce_(self, *tensors):
        """In-place sum over ranks.  Several small tensors are flattened into ONE call (latency-bound)."""
        if self.world == 1:
            return
        self.n_allreduce += 1
        self.bytes_allreduce += sum(t.numel() * t.element_size() for t in tensors)
        if len(tensors) == 1:
            self.dist.all_reduce(tensors[0], group=self.group)
            return
        flat = torch.cat([t.reshape(-1) for t in tensors])
        self.dist.all_reduce(flat, group=self.group)
        off = 0
        for t in tensors:
            t.copy_(flat[off:off + t.numel()].view_as(t))
            off += t.numel()

    def sum_scalar(self, value, device):
        t = torch.tensor([float(value)], dtype=torch.float64, device=device)
        self.all_reduce_(t)
        return float(t[0])


class ShardDataset:
    """This rank's row block, with the reference's dataset-iterator interface
    (src/xGPR/data_handling/online_data_handling.py:54-94, data_handling_baseclass.py:40-75).

    x / y / sequence_lengths may be NumPy arrays, pinned host tensors or CUDA tensors (a resident shard
    avoids the per-pass H2D copies).  ``trainy_mean`` / ``trainy_std`` and ``global_ndatapoints`` are
    GLOBAL statistics (use ``build_shard_dataset``)."""

    def __init__(self, xdata, ydata=None, sequence_lengths=None, chunk_size=2000, trainy_mean=0.0,
                 trainy_std=1.0, global_ndatapoints=None):
        self._xdata, self._ydata, self._sequence_lengths = xdata, ydata, sequence_lengths
        self._chunk_size = int(chunk_size)
        self._trainy_mean, self._trainy_std = float(trainy_mean), float(trainy_std)
        self._n_local = int(xdata.shape[0])
        self._n_global = int(global_ndatapoints) if global_ndatapoints is not None else self._n_local

    def get_chunk_size(self):
        return self._chunk_size

    def get_ndatapoints(self):
        return self._n_global

    def get_local_ndatapoints(self):
        return self._n_local

    def get_xdim(self):
        return tuple(self._xdata.shape)

    def get_ymean(self):
        return self._trainy_mean

    def get_ystd(self):
        return self._trainy_std

    def _ychunk(self, lo, hi):
        y = self._ydata[lo:hi]
        if isinstance(y, torch.Tensor):
            return (y.to(torch.float64) - self._trainy_mean) / self._trainy_std
        y = y.astype(np.float64)
        y -= self._trainy_mean
        y /= self._trainy_std
        return y

    def get_chunked_data(self):
        for lo in range(0, self._n_local, self._chunk_size):
            hi = min(lo + self._chunk_size, self._n_local)
            sl = None if self._sequence_lengths is None else self._sequence_lengths[lo:hi]
            yield self._xdata[lo:hi, ...], self._ychunk(lo, hi), sl

    def get_chunked_x_data(self):
        for lo in range(0, self._n_local, self._chunk_size):
            hi = min(lo + self._chunk_size, self._n_local)
            sl = None if self._sequence_lengths is None else self._sequence_lengths[lo:hi]
            yield self._xdata[lo:hi, ...], sl

    def get_coalesced_data(self, min_rows, with_y=True):
        """The same rows in the same order as get_chunked_data / get_chunked_x_data, in work chunks of a whole number of
        dataset chunks totalling >= min_rows rows.  The reference's default chunk_size is 2,000 rows
        (online_data_handling.py:26) -- 43 us of GPU work at 46 M rows/s, behind three launches and the Python of a chunk
        iteration -- so the passes whose result is a sum over rows (Z^T y, Z^T Z, the CG matvec, the Nystrom sketch) walk
        the in-memory shard in larger slices; only fp64 summation order changes.  Passes that draw per-chunk random
        subsamples (the NMLL gradient, the rank-ratio probe) keep the dataset's own chunking."""
        step = self._chunk_size * max(1, int(min_rows) // self._chunk_size)
        for lo in range(0, self._n_local, step):
            hi = min(lo + step, self._n_local)
            sl = None if self._sequence_lengths is None else self._sequence_lengths[lo:hi]
            if with_y:
                yield self._xdata[lo:hi, ...], self._ychunk(lo, hi), sl
            else:
                yield self._xdata[lo:hi, ...], sl


def build_shard_dataset(xdata, ydata, sequence_lengths=None, chunk_size=2000, comm=None):
    """y is standardised with the GLOBAL mean / std (dataset_builder.py:14-70 computes them over the whole
    training set): one all-reduce of (n, sum y, sum y^2)."""
    comm = comm or RowShardComm()
    yv = ydata.detach().double().cpu().numpy() if isinstance(ydata, torch.Tensor) else np.asarray(ydata, np.float64)
    dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    n = comm.sum_scalar(yv.shape[0], dev)
    mean = comm.sum_scalar(yv.sum(), dev) / n
    var = comm.sum_scalar(((yv - mean) ** 2).sum(), dev) / n       # np.std: population std
    return ShardDataset(xdata, ydata, sequence_lengths, chunk_size, mean, float(np.sqrt(var)), int(round(n)))


# ----------------------------------------------------------------------------------------- accumulation passes
def _new(shape, like_device):
    return torch.zeros(shape, dtype=torch.float64, device=like_device)


def _work_rows(kernel, budget_bytes=4 << 30, cap=65536):
    """Rows per work chunk: about `budget_bytes` of fp64 features (8 M bytes per row), at most `cap`."""
    return int(max(1, min(cap, budget_bytes // (8 * max(1, kernel.get_num_rffs())))))


def _xy_chunks(dataset, kernel):
    fn = getattr(dataset, "get_coalesced_data", None)
    return fn(_work_rows(kernel), True) if fn is not None else dataset.get_chunked_data()


def _x_chunks(dataset, kernel):
    fn = getattr(dataset, "get_coalesced_data", None)
    return fn(_work_rows(kernel), False) if fn is not None else dataset.get_chunked_x_data()


def _transformed_chunks(dataset, kernel):
    """kernel.transform_x over the shard's chunks; on the GPU through the kernel's double-buffered transform_stream, so
    a host-resident chunk's copy runs under the previous chunk's feature generation."""
    stream = getattr(kernel, "transform_stream", None)
    if stream is not None and kernel.device == "cuda":
        yield from stream(_x_chunks(dataset, kernel))
        return
    for xdata, ldata in _x_chunks(dataset, kernel):
        yield kernel.transform_x(xdata, ldata)


def _transformed_xy_chunks(dataset, kernel):
    """(Z, y) per chunk, as kernel.transform_x_y over dataset.get_chunked_data(), through the double-buffered stream."""
    stream = getattr(kernel, "transform_stream", None)
    if stream is None or kernel.device != "cuda":
        for xin, yin, ldata in _xy_chunks(dataset, kernel):
            yield kernel.transform_x_y(xin, yin, ldata)
        return
    from .kernels.base import to_device
    ys = []

    def x_chunks():
        for xin, yin, ldata in _xy_chunks(dataset, kernel):
            ys.append(yin)
            yield xin, ldata

    for z in stream(x_chunks()):
        yin = ys.pop(0)
        if isinstance(yin, np.ndarray):
            yin = yin.astype(np.float64)
        yield z, to_device(yin, torch.float64)


def _as_f64(y):
    return y.astype(np.float64) if isinstance(y, np.ndarray) else y


def _available_hbm(dev):
    """Bytes a new allocation can use: free device memory plus what torch's caching allocator holds but is not using."""
    free, _ = torch.cuda.mem_get_info(dev)
    return free + torch.cuda.memory_reserved(dev) - torch.cuda.memory_allocated(dev)


def _device_of(kernel):
    return torch.device("cuda", torch.cuda.current_device()) if kernel.device == "cuda" else torch.device("cpu")


def _fused_ok(kernel, ops):
    """Fused featurise->matvec is offered by the fixed-vector SORF kernels on the CUDA path (<= 1024 inputs)."""
    return ops is _cuda_ops and getattr(kernel, "supports_fused_matvec", False) and kernel.get_xdim()[-1] <= 1024


def calc_zty(dataset, kernel, comm=None, ops=_cuda_ops):
    """z^T y and y^T y over the whole (sharded) dataset.  exact_nmll_calcs.py:13-38."""
    comm = comm or RowShardComm()
    dev = _device_of(kernel)
    z_trans_y, y_trans_y = _new((kernel.get_num_rffs(),), dev), _new((1,), dev)
    if _fused_ok(kernel, ops):
        # fused featurise -> Z^T y: the features are regenerated in registers and never written (rbf_ztv); the next
        # chunk's x and y are copied under this chunk's kernel (stage_stream)
        def chunks():
            for xin, yin, ldata in _xy_chunks(dataset, kernel):
                yield xin, ldata, yin

        for xin, _, (y,) in kernel.stage_stream(chunks(), extra=lambda item: (_as_f64(item[2]),)):
            y = y.to(torch.float64).contiguous()
            kernel.fused_ztv(kernel.scale_staged(xin), y, z_trans_y)
            y_trans_y += y @ y
        comm.all_reduce_(z_trans_y, y_trans_y)
        return z_trans_y, float(y_trans_y[0])
    for z, y in _transformed_xy_chunks(dataset, kernel):
        ops.zty_accum(z, y, z_trans_y)
        y_trans_y += y @ y
    comm.all_reduce_(z_trans_y, y_trans_y)
    return z_trans_y, float(y_trans_y[0])


def _gram_accumulator(ops, gram, gram_slices):
    """Returns (accumulate(z, ztz), finish(ztz)).  gram="tc": tensor-core split-int8 Gram (tcgen05, exact int32
    accumulation of `gram_slices` = (planes, bits per digit) int8 digit planes -- default 5 x 8 bits, error ~4e-13 of the
    column scale per row; an int means that many 7-bit planes -- csrc/gram_tc_kernels.cu), upper triangle per chunk and one mirror at the
    end; gram="f64": the DFMA kernel.  Providers without the tensor-core entry point (the CPU test provider) use f64."""
    if gram == "tc" and hasattr(ops, "ztz_accum_tc"):
        ns, bits = gram_slices if isinstance(gram_slices, tuple) else (gram_slices, 7)
        return (lambda z, ztz: ops.ztz_accum_tc(z, ztz, num_slices=ns, mirror=False, digit_bits=bits)), \
            ops.symmetrize_upper
    if gram not in ("tc", "f64"):
        raise ValueError("gram must be 'tc' or 'f64'")
    return ops.ztz_accum, (lambda ztz: None)


def calc_design_mat(dataset, kernel, comm=None, ops=_cuda_ops, gram="tc", gram_slices=(5, 8)):
    """z^T z, z^T y, y^T y.  exact_nmll_calcs.py:42-78; one all-reduce of the M x M Gram per pass."""
    comm = comm or RowShardComm()
    dev = _device_of(kernel)
    m = kernel.get_num_rffs()
    z_trans_z, z_trans_y, y_trans_y = _new((m, m), dev), _new((m,), dev), _new((1,), dev)
    accumulate, finish = _gram_accumulator(ops, gram, gram_slices)
    for z, y in _transformed_xy_chunks(dataset, kernel):
        ops.zty_accum(z, y, z_trans_y)
        accumulate(z, z_trans_z)
        y_trans_y += y @ y
    finish(z_trans_z)
    comm.all_reduce_(z_trans_z)
    comm.all_reduce_(z_trans_y, y_trans_y)
    return z_trans_z, z_trans_y, float(y_trans_y[0])


def direct_weight_calc(chol_z_trans_z, z_trans_y, kernel):
    """Cholesky + two triangular solves; adds lambda^2 to the diagonal.  exact_nmll_calcs.py:82-112."""
    lambda_p = kernel.get_hyperparams(logspace=False)[0]
    chol_z_trans_z.diagonal().add_(lambda_p ** 2)
    chol = torch.linalg.cholesky(chol_z_trans_z)
    w = torch.linalg.solve_triangular(chol, z_trans_y[:, None], upper=False)
    w = torch.linalg.solve_triangular(chol.T, w, upper=True)
    return chol, w[:, 0]


def calc_weights_exact(dataset, kernel, comm=None, ops=_cuda_ops, gram="tc", gram_slices=(5, 8)):
    """exact_fitting_toolkit.py:16-36.  lambda^2 is added here AND inside direct_weight_calc, exactly as
    the reference does: the exact fit solves (Z^T Z + 2 lambda^2 I) w = Z^T y."""
    z_trans_z, z_trans_y, _ = calc_design_mat(dataset, kernel, comm, ops, gram, gram_slices)
    lambda_p = kernel.get_hyperparams(logspace=False)[0]
    z_trans_z.diagonal().add_(lambda_p ** 2)
    _, weights = direct_weight_calc(z_trans_z, z_trans_y, kernel)
    return weights, 1, []


def calc_var_design_mat(dataset, kernel, variance_rffs, comm=None, ops=_cuda_ops, gram="tc", gram_slices=(5, 8)):
    """exact_nmll_calcs.py:116-139: Gram of the first variance_rffs columns."""
    comm = comm or RowShardComm()
    z_trans_z = _new((variance_rffs, variance_rffs), _device_of(kernel))
    accumulate, finish = _gram_accumulator(ops, gram, gram_slices)
    for z in _transformed_chunks(dataset, kernel):
        accumulate(z[:, :variance_rffs].contiguous(), z_trans_z)
    finish(z_trans_z)
    comm.all_reduce_(z_trans_z)
    return z_trans_z


def calc_variance_exact(kernel, dataset, variance_rffs, comm=None, ops=_cuda_ops):
    """exact_fitting_toolkit.py:40-68."""
    z_trans_z = calc_var_design_mat(dataset, kernel, variance_rffs, comm, ops)
    z_trans_z.diagonal().add_(kernel.get_lambda() ** 2)
    return torch.linalg.pinv(z_trans_z)


# ----------------------------------------------------------------------------------------- conjugate gradients
_TOKENS = iter(range(1, 1 << 62))


def _token(obj):
    """A process-unique token stored ON the object: unlike id(), it is never reused for a different object after the
    first one is garbage collected (a tuning loop rebuilding datasets / kernels would otherwise be served stale features)."""
    tok = getattr(obj, "_xb_token", None)
    if tok is None:
        tok = next(_TOKENS)
        try:
            obj._xb_token = tok
        except AttributeError:              # objects without a __dict__: fall back to a weak identity
            return ("id", id(obj))
    return tok


class ConjugateGrad:
    """Batched-RHS preconditioned CG on (Z^T Z + lambda^2 I).  cg_tools.py:15-156 (GPU_ConjugateGrad)."""

    def __init__(self, comm=None, ops=_cuda_ops):
        self.comm = comm or RowShardComm()
        self.ops = ops
        self._scratch = None
        self._staged = None
        self._resident = None
        self._resident_c = None
        self._encoded = None
        self.keep_resident = True
        # several right-hand sides: stream the float32 encoding of Z (4 bytes per feature) instead of fp64 Z
        self.encoded_multi_rhs = os.environ.get("XGPR_B200_MULTIRHS_F64", "0") == "0"
        self.encoded_fraction = float(os.environ.get("XGPR_B200_ENCODED_FRACTION", "0.85"))   # share of the free HBM the kept chunks of the encoding may take
        self.one_pass = True
        # one right-hand side on the GPU: scalar recurrences on the device, no per-iteration host sync (_fit_device)
        self.device_cg = os.environ.get("XGPR_B200_DEVICE_CG", "1") != "0"
        self.use_graph = os.environ.get("XGPR_B200_CG_GRAPH", "1") != "0"
        self.poll_lag = 2
        self.last_fit_info = {}
        self._prep_event = None

    def _staged_blocks(self, dataset, kernel, budget_bytes=8 << 30):
        """The sigma-scaled device copies of the shard's x blocks for the fused passes.  x * sigma does not change during
        a fit, so the copies (4 * D bytes per row; 256 MB for cfg2's 1M x 64) are staged ONCE and reused by every CG
        iteration -- with a device-resident shard as one block, i.e. two launches per iteration -- unless they would
        exceed `budget_bytes`, in which case each chunk is staged per pass as before."""
        key = (_token(dataset), _token(kernel), float(kernel.get_hyperparams(logspace=False)[1]))
        if self._staged is not None and self._staged[0] == key:
            return self._staged[1]
        xdata = dataset._xdata
        nbytes = int(np.prod(xdata.shape)) * (8 if kernel.double_precision else 4)
        if nbytes > budget_bytes:
            return ((kernel.stage_scaled(x), torch.empty(x.shape[0], dtype=torch.float64, device=_device_of(kernel)))
                    for x, _ in _x_chunks(dataset, kernel))
        if isinstance(xdata, torch.Tensor) and xdata.is_cuda:
            pieces = [xdata]
        else:
            pieces = [x for x, _ in _x_chunks(dataset, kernel)]
        blocks = []
        for x in pieces:
            xs = kernel.stage_scaled(x)
            blocks.append((xs, torch.empty(xs.shape[0], dtype=torch.float64, device=xs.device)))
        self._staged = (key, blocks)
        return blocks

    def _matvec(self, dataset, kernel, vec, matvec):
        """matvec <- sum_chunks Z^T (Z vec) + lambda^2 vec.  cg_tools.py:26-53.  The per-rank partial sum is
        all-reduced ONCE per call (one NCCL call per CG iteration); lambda^2 vec is added after it."""
        matvec.zero_()
        vec = vec.contiguous()
        if _fused_ok(kernel, self.ops) and vec.shape[1] == 1:
            # one right-hand side: two fused feature passes per block, Z never written to HBM
            v1, w1 = vec[:, 0].contiguous(), matvec[:, 0]
            resident = self._resident_encoding(dataset, kernel)
            if resident is not None:             # one pass over the float32 encoding of Z kept in HBM
                kernel.resident_zzv(resident, v1, w1)
                self.comm.all_reduce_(matvec)
                matvec.add_(vec, alpha=kernel.get_lambda() ** 2)
                return
            one_pass = self.one_pass and getattr(kernel, "fused_zzv_ok", lambda: False)()
            for xs, t in self._staged_blocks(dataset, kernel):
                if one_pass:                     # t = z . v reduced on chip: one SORF per frequency per iteration
                    kernel.fused_zzv(xs, v1, w1)
                    continue
                t.zero_()
                kernel.fused_zv(xs, v1, t)
                kernel.fused_ztv(xs, t, w1)
            self.comm.all_reduce_(matvec)
            matvec.add_(vec, alpha=kernel.get_lambda() ** 2)
            return
        encoded = (self.encoded_multi_rhs and self.ops is _cuda_ops and kernel.device == "cuda"
                   and getattr(kernel, "encoding_ok", lambda: False)())
        if encoded:
            # Z = double(C) * scale exactly (column 0 = 1 with an intercept): C v_s with v_s = diag(s) v, then diag(s) C^T t
            scale = kernel.feature_scale()
            vs = vec * scale
            if kernel.fit_intercept:
                vs[0] = vec[0]
        for z in (self._encoded_chunks(dataset, kernel) if encoded else self._feature_chunks(dataset, kernel)):
            if self._scratch is None or self._scratch.shape[0] < z.shape[0] or self._scratch.shape[1] != vec.shape[1] \
                    or self._scratch.device != z.device:
                self._scratch = torch.empty((z.shape[0], vec.shape[1]), dtype=torch.float64, device=z.device)
            if encoded:
                self.ops.cg_matvec_accum_c32(z, vs, matvec, scale, kernel.fit_intercept, self._scratch)
            else:
                self.ops.cg_matvec_accum(z, vec, matvec, self._scratch)
        self.comm.all_reduce_(matvec)
        matvec.add_(vec, alpha=kernel.get_lambda() ** 2)

    def _resident_encoding(self, dataset, kernel, fraction=0.6, piece=65536):
        """The shard's features as ONE resident float32 matrix C (Z = double(C) * scale exactly; 4 M bytes per row: cfg2's
        1M x 16,384 = 65.5 GB), built once per solve when the kernel supports it and C plus one fp64 staging piece fit in
        `fraction` of the free HBM; None otherwise (the fused regenerate-per-iteration kernels are used instead)."""
        if not self.keep_resident or not getattr(kernel, "resident_ok", lambda: False)():
            return None
        key = (_token(dataset), _token(kernel), tuple(np.asarray(kernel.get_hyperparams(logspace=False)).tolist()))
        if self._resident_c is not None and self._resident_c[0] == key:
            return self._resident_c[1]
        self._resident_c = None
        dev = _device_of(kernel)
        rows, m = dataset.get_local_ndatapoints(), kernel.get_num_rffs()
        free = _available_hbm(dev)
        if rows * m * 4 + min(piece, rows) * m * 8 > fraction * free:
            self._resident_c = (key, None)
            return None
        try:
            c = torch.empty((rows, m), dtype=torch.float32, device=dev)
        except torch.cuda.OutOfMemoryError:      # fragmented pool: keep the fused regenerate-per-iteration path
            self._resident_c = (key, None)
            return None
        lo = 0
        direct = getattr(kernel, "resident_generate", None) if kernel.get_xdim()[-1] <= 1024 else None
        for xdata, ldata in _x_chunks(dataset, kernel):
            if direct is not None:               # the feature kernel writes the float32 values itself
                direct(xdata, c[lo:lo + xdata.shape[0]])
                lo += xdata.shape[0]
                continue
            for p0 in range(0, xdata.shape[0], piece):
                z = kernel.transform_x(xdata[p0:p0 + piece], None if ldata is None else ldata[p0:p0 + piece])
                kernel.resident_encode(z, c[lo:lo + z.shape[0]])
                lo += z.shape[0]
                del z
        if kernel.fit_intercept:
            # column 0 stands for the intercept: the one-pass matvec substitutes 1 for it on the fly, the other readers of
            # the encoding (the preconditioner's accumulation pass) expect the stored value to be 1.0f
            c[:, 0] = 1.0
        self._resident_c = (key, c)
        return c

    def _feature_chunks(self, dataset, kernel, fraction=0.5):
        """The shard's feature chunks for the materialised (multi-RHS) matvec.  Z does not change during a solve, so when
        the whole shard's Z (8 M bytes per row) fits in `fraction` of the free HBM it is generated ONCE and kept resident
        for every iteration (cfg5 on 8 GPUs: 250k rows x 32,768 RFFs = 65 GB of 180 GB); otherwise it is regenerated per
        iteration, as the reference does (cg_tools.py:42-44)."""
        key = (_token(dataset), _token(kernel), tuple(np.asarray(kernel.get_hyperparams(logspace=False)).tolist()))
        if self._resident is not None and self._resident[0] == key:
            yield from self._resident[1]
            return
        self._resident = None
        dev = _device_of(kernel)
        keep = None
        if self.keep_resident and dev.type == "cuda":
            rows = dataset.get_local_ndatapoints()
            if rows * kernel.get_num_rffs() * 8 <= fraction * _available_hbm(dev):
                keep = []
        for z in _transformed_chunks(dataset, kernel):
            if keep is not None:
                keep.append(z)
            yield z
        if keep is not None:
            self._resident = (key, keep)

    def _encoded_chunks(self, dataset, kernel):
        """The shard's features for the multi-RHS matvec as chunks of the float32 encoding C (4 M bytes per row, half of
        fp64 Z; lossless for the float32-FHT kernels).  C does not change during a solve: the first pass keeps every
        chunk it can within `fraction` of the free HBM (cfg5 on one GPU: 2M x 32,768 = 262 GB, so about half the rows
        stay; on 8 GPUs all of them), the others are regenerated per iteration into one reusable buffer."""
        key = (_token(dataset), _token(kernel), tuple(np.asarray(kernel.get_hyperparams(logspace=False)).tolist()))
        if self._encoded is None or self._encoded[0] != key:
            self._encoded = None
            dev = _device_of(kernel)
            row_bytes = kernel.get_num_rffs() * 4
            work_bytes = min(_work_rows(kernel), dataset.get_local_ndatapoints()) * row_bytes
            budget = self.encoded_fraction * _available_hbm(dev) - work_bytes if self.keep_resident else 0
            self._encoded = (key, {"kept": {}, "budget": budget, "work": None})
        state = self._encoded[1]
        m = kernel.get_num_rffs()
        for idx, (xdata, _) in enumerate(_x_chunks(dataset, kernel)):
            c = state["kept"].get(idx)
            if c is not None:
                yield c
                continue
            rows = xdata.shape[0]
            out = None
            if state["budget"] >= rows * m * 4:
                try:
                    out = torch.empty((rows, m), dtype=torch.float32, device=_device_of(kernel))
                    state["kept"][idx] = out
                    state["budget"] -= rows * m * 4
                except torch.cuda.OutOfMemoryError:              # fragmented pool: stop keeping chunks
                    state["budget"] = 0
            if out is None:
                work = state["work"]
                if work is None or work.shape[0] < rows:
                    state["work"] = work = torch.empty((rows, m), dtype=torch.float32, device=_device_of(kernel))
                out = work[:rows]
            kernel.resident_generate(xdata, out)
            if kernel.fit_intercept:
                out[:, 0] = 1.0
            yield out

    def resident_for_preconditioner(self, dataset, kernel):
        """The shard's resident float32 encoding, built NOW on the current stream, when the solve that follows will use it
        (same conditions as prepare_async) -- so that the preconditioner's accumulation pass can read it instead of
        generating fp64 features a second time; None otherwise.  XGPR_B200_PRECOND_F64=1 keeps the fp64 pass."""
        if os.environ.get("XGPR_B200_PRECOND_F64", "0") != "0":
            return None
        if not (self.device_cg and self.ops is _cuda_ops and kernel.device == "cuda" and _fused_ok(kernel, self.ops)
                and self.keep_resident and getattr(kernel, "resident_ok", lambda: False)()):
            return None
        return self._resident_encoding(dataset, kernel)

    def prepare_async(self, dataset, kernel):
        """Start building the resident float32 encoding of the shard's features on a SIDE stream and return at once.
        Meant to be called while the GPU is under-used -- `RandNysPreconditioner(..., overlap_hook=...)` calls it between
        its accumulation pass and its serial factorisation (a few ms of small cuSOLVER kernels, replicated on every rank:
        the Amdahl term of the sharded fit), so the solve that follows finds its features ready.  No-op when the solve
        would not use the resident encoding.  `fit` waits for the side stream before its first matvec."""
        if not (self.device_cg and self.ops is _cuda_ops and kernel.device == "cuda" and _fused_ok(kernel, self.ops)
                and self.keep_resident and getattr(kernel, "resident_ok", lambda: False)()):
            return
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            c = self._resident_encoding(dataset, kernel)
            done = torch.cuda.Event()
            done.record(side)
        if c is not None:
            c.record_stream(cur)
        self._prep_event = done

    # ---- one right-hand side, device-resident recurrences -----------------------------------------------------------
    def _device_cg_ok(self, kernel, resid, nmll_settings):
        return (self.device_cg and self.ops is _cuda_ops and not nmll_settings and resid.shape[2] == 1 and resid.is_cuda
                and kernel.device == "cuda" and _fused_ok(kernel, self.ops))

    def _fit_device(self, dataset, kernel, preconditioner, resid, maxiter, tol, verbose):
        """cg_tools.py:57-156 for one right-hand side with NO host round trip inside the loop.  Per iteration, all on the
        current stream: [memset w, matvec kernel(s), ONE all-reduce of w, cg_update_a, two preconditioner GEMVs (+ one
        scaling), cg_update_b] -- alpha, beta, the residual history and the convergence flag stay in a device `state`
        (csrc/cg_kernels.cu).  After one eager iteration (cuBLAS / NCCL / kernel attributes initialised) the iteration is
        captured ONCE in a CUDA graph and replayed.  The host reads the convergence flag with a fixed lag of `poll_lag`
        iterations (8-byte async copies into pinned memory), so the GPU always has the next iteration queued; iterations
        launched after convergence are no-ops on the device (the resident matvec is gated on the flag as well), hence x, the
        iteration count and the loss history are exactly those of stopping at the right iteration.  With several ranks
        every rank sees bit-identical all-reduced w, takes the same decision at the same iteration, and so issues the same
        number of collectives."""
        ops = self.ops
        if self._prep_event is not None:       # prepare_async built the resident encoding on a side stream
            torch.cuda.current_stream().wait_event(self._prep_event)
            self._prep_event = None
        dev, m = resid.device, resid.shape[0]
        lam2 = float(kernel.get_lambda()) ** 2
        r = resid[:, 0, 0].clone().contiguous()
        x, w = torch.zeros(m, dtype=torch.float64, device=dev), torch.zeros(m, dtype=torch.float64, device=dev)
        z, p, ps = torch.empty_like(r), torch.empty_like(r), torch.empty_like(r)
        state = torch.zeros(ops.CG_STATE_LEN, dtype=torch.float64, device=dev)
        hist = torch.zeros(maxiter, dtype=torch.float64, device=dev)
        skip = state[5:6]
        if preconditioner is not None:
            u_mat = preconditioner.u_mat
            coef = (preconditioner.inv_eig * preconditioner.prefactor - 1.0).contiguous()
            tsm = torch.empty(u_mat.shape[1], dtype=torch.float64, device=dev)

        def apply_precond():                 # z = P^-1 r = r + U ((inv_eig * prefactor - 1) o (U^T r)); batch_matvec
            if preconditioner is None:
                z.copy_(r)
                return
            torch.mv(u_mat.T, r, out=tsm)
            tsm.mul_(coef)
            torch.addmv(r, u_mat, tsm, out=z)

        resident = self._resident_encoding(dataset, kernel)
        scale, icpt = kernel.feature_scale(), kernel.fit_intercept
        one_pass = self.one_pass and getattr(kernel, "fused_zzv_ok", lambda: False)()
        blocks = None if resident is not None else list(self._staged_blocks(dataset, kernel))

        def iteration():
            w.zero_()
            if resident is not None:
                kernel.resident_zzv_scaled(resident, ps, w, skip)
            else:
                for xs, t in blocks:
                    if one_pass:
                        kernel.fused_zzv(xs, p, w)
                    else:
                        t.zero_()
                        kernel.fused_zv(xs, p, t)
                        kernel.fused_ztv(xs, t, w)
            self.comm.all_reduce_(w)
            ops.cg_update_a(w, p, x, r, lam2, state, hist)
            apply_precond()
            ops.cg_update_b(r, z, p, state, ps, scale, icpt)

        apply_precond()
        p.copy_(z)
        torch.mul(p, scale, out=ps)
        if icpt:
            ps[0:1].copy_(p[0:1])
        state[0:1].copy_((r * z).sum().reshape(1))
        state[1:2].copy_(torch.linalg.norm(r).reshape(1))
        state[2:3].fill_(float(tol))

        n_ar0, b_ar0 = self.comm.n_allreduce, self.comm.bytes_allreduce
        lag = max(1, int(self.poll_lag))
        flags = torch.zeros(lag + 2, dtype=torch.float64).pin_memory()
        events = [torch.cuda.Event() for _ in range(lag + 2)]
        graph, launched, graphed = None, 0, False
        for k in range(maxiter):
            if k >= lag:
                slot = (k - lag) % (lag + 2)
                events[slot].synchronize()
                if flags[slot] != 0.0:
                    break
            if k == 1 and self.use_graph and graph is None:
                graph = self._capture(iteration)
                graphed = graph is not None
            if graph is not None and k >= 1:
                graph.replay()
            else:
                iteration()
            slot = k % (lag + 2)
            flags[slot:slot + 1].copy_(skip, non_blocking=True)
            events[slot].record()
            launched += 1
        if self.comm.world > 1:               # graph replays issue the captured all-reduce without passing through comm
            self.comm.n_allreduce = n_ar0 + launched
            self.comm.bytes_allreduce = b_ar0 + launched * w.numel() * w.element_size()
        st = state.cpu()
        done = bool(st[5] != 0.0)
        niter = int(st[6]) if done else launched
        losses = hist[:niter].cpu().tolist()
        self.last_fit_info = {"device_cg": True, "cuda_graph": graphed, "iterations_launched": launched,
                              "resident": resident is not None}
        if verbose:
            print(f"{niter} iterations complete.")
        return x, done, niter, losses

    @staticmethod
    def _capture(iteration):
        """One CG iteration as a CUDA graph (NCCL's all-reduce is capturable); None if capture is not possible.
        Uses CUDAGraph.capture_begin / capture_end directly: the `torch.cuda.graph` context manager also runs
        gc.collect(), a device-wide synchronize and empty_cache() -- milliseconds of serial host work (and a cold allocator
        afterwards) that a 0.1 s sharded solve cannot afford."""
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        graph = torch.cuda.CUDAGraph()
        began = False
        try:
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                graph.capture_begin(capture_error_mode="thread_local")
                began = True
                iteration()
                graph.capture_end()
                began = False
            cur.wait_stream(side)
            return graph
        except Exception as exc:             # keep the stream-ordered eager launches
            if began:
                try:
                    with torch.cuda.stream(side):
                        graph.capture_end()
                except Exception:
                    pass
            warnings.warn(f"CUDA-graph capture of the CG iteration failed ({exc}); launching eagerly.")
            return None

    def fit(self, dataset, kernel, preconditioner, resid, maxiter=200, tol=1e-4, verbose=True,
            nmll_settings=False):
        if self._device_cg_ok(kernel, resid, nmll_settings):
            return self._fit_device(dataset, kernel, preconditioner, resid, maxiter, tol, verbose)
        self.last_fit_info = {"device_cg": False}
        converged = False
        target = resid[:, 0, :].clone()
        init_norms = torch.linalg.norm(target, dim=0)
        dev, nr, nrhs = resid.device, resid.shape[0], resid.shape[2]
        z_k = torch.zeros((nr, 2, nrhs), dtype=torch.float64, device=dev)
        p_k = torch.zeros((nr, 2, nrhs), dtype=torch.float64, device=dev)
        alphas, betas, losses = [], [], []
        x_k = torch.zeros((nr, nrhs), dtype=torch.float64, device=dev)
        w = torch.zeros((nr, nrhs), dtype=torch.float64, device=dev)

        z_k[:, 0, :] = resid[:, 0, :] if preconditioner is None else preconditioner.batch_matvec(resid[:, 0, :])
        p_k[:, 0, :] = z_k[:, 0, :]
        next_col, current_col = 1, 0
        niter = 0
        for niter in range(maxiter):
            self._matvec(dataset, kernel, p_k[:, current_col, :], w)
            alpha = (resid[:, current_col, :] * z_k[:, current_col, :]).sum(dim=0) / \
                    (p_k[:, current_col, :] * w).sum(dim=0)
            x_k += alpha[None, :] * p_k[:, current_col, :]
            resid[:, next_col, :] = resid[:, current_col, :] - alpha[None, :] * w
            # convergence is judged on the PREVIOUS residual, as in the reference (cg_tools.py:116,139)
            err = torch.linalg.norm(resid[:, current_col, :], dim=0) / init_norms
            if preconditioner is None:
                z_k[:, next_col, :] = resid[:, next_col, :]
            else:
                z_k[:, next_col, :] = preconditioner.batch_matvec(resid[:, next_col, :])
            beta = (resid[:, next_col, :] * z_k[:, next_col, :]).sum(dim=0) / \
                   (resid[:, current_col, :] * z_k[:, current_col, :]).sum(dim=0)
            p_k[:, next_col, :] = z_k[:, next_col, :] + beta[None, :] * p_k[:, current_col, :]
            err_host = err.tolist()                      # the iteration's only host sync
            if nmll_settings:
                alphas.append(alpha.clone())
                betas.append(beta.clone())
            else:
                losses.append(err_host[0])
            next_col, current_col = abs(next_col - 1), abs(current_col - 1)
            if niter % 5 == 0 and verbose:
                print(f"{niter} iterations complete.")
            if max(err_host) < tol:
                converged = True
                break

        if nmll_settings:
            alphas, betas = torch.stack(alphas), torch.stack(betas)
            return x_k, alphas[:, 1:].cpu().numpy(), betas[:, 1:].cpu().numpy()
        if x_k.shape[1] > 1:
            return x_k, converged, niter + 1, losses
        return x_k[:, 0], converged, niter + 1, losses


def cg_fit_lib_internal(kernel, dataset, cg_tol=1e-4, max_iter=500, preconditioner=None, verbose=True,
                        comm=None, ops=_cuda_ops, cg_operator=None):
    """cg_fitting_toolkit.py:18-70: rhs = Z^T y / N, zero start, result * N."""
    comm = comm or RowShardComm()
    cg_operator = cg_operator or ConjugateGrad(comm, ops)
    if preconditioner is None:
        z_trans_y, _ = calc_zty(dataset, kernel, comm, ops)
    else:
        z_trans_y = preconditioner.get_zty()
    resid = torch.zeros((kernel.get_num_rffs(), 2, 1), dtype=torch.float64, device=z_trans_y.device)
    resid[:, 0, :] = z_trans_y[:, None] / dataset.get_ndatapoints()
    weights, converged, n_iter, losses = cg_operator.fit(dataset, kernel, preconditioner, resid, max_iter, cg_tol,
                                                         verbose, nmll_settings=False)
    weights *= dataset.get_ndatapoints()
    if not converged:
        warnings.warn("Conjugate gradients failed to converge! Try refitting the model with updated settings.")
    if verbose:
        print(f"CG iterations: {n_iter}")
    return weights, n_iter, losses


def fit_cg(kernel, dataset, max_rank=512, method="srht", cg_tol=1e-6, max_iter=500, random_state=123, verbose=False,
           comm=None, ops=_cuda_ops):
    """`build_preconditioner` + `fit(mode="cg", preconditioner=...)` of the reference model class
    (model_baseclass.py:225-260, xgp_regression.py:381-470) as one call: the preconditioner build's serial factorisation
    overlaps with the construction of the solve's resident features.  Returns (weights, n_iter, losses, preconditioner)."""
    comm = comm or RowShardComm()
    op = ConjugateGrad(comm, ops)
    pre = RandNysPreconditioner(kernel, dataset, max_rank, verbose, random_state, method, comm, ops,
                                overlap_hook=lambda: op.prepare_async(dataset, kernel),
                                encoded_source=lambda: op.resident_for_preconditioner(dataset, kernel))
    weights, n_iter, losses = cg_fit_lib_internal(kernel, dataset, cg_tol, max_iter, pre, verbose, comm, ops, op)
    return weights, n_iter, losses, pre


# ----------------------------------------------------------------------------------------- randomised Nystrom
def _single_pass_srht_zty(dataset, kernel, compressor, acc_results, z_trans_y, ops):
    """rand_nys_constructors.py:96-123: acc (rank, M) += SRHT(Z)[:, perm[:rank]]^T Z ; z^T y ; y^T y.
    The sketch product runs on the tensor cores (exact-int8 planes, fp64-level error) when the provider has it."""
    y_trans_y = _new((1,), acc_results.device)
    atb = getattr(ops, "atb_accum_tc", None) or ops.atb_accum
    for z, y in _transformed_xy_chunks(dataset, kernel):
        ops.zty_accum(z, y, z_trans_y)
        y_trans_y += y @ y
        atb(compressor.transform_x(z), z, acc_results)
    return y_trans_y


def _single_pass_srht_zty_encoded(dataset, kernel, compressor, acc_results, z_trans_y, ops, c_mat):
    """_single_pass_srht_zty fed from the shard's resident float32 encoding C (rows in dataset order; Z = double(C) * scale
    exactly): no second featurise, and Z^T y, the SRHT sample and the digit-plane passes of the tensor-core sketch read 4
    bytes per feature.  Returns None when a step is not available for this shape (the caller falls back to fp64 Z)."""
    scale, icpt = kernel.feature_scale(), bool(kernel.fit_intercept)
    if not hasattr(compressor, "transform_encoded") or not hasattr(ops, "atb_accum_tc_c32") or c_mat.shape[1] % 4:
        return None
    if compressor.transform_encoded(c_mat[:1], scale, icpt) is None:
        return None
    y_trans_y = _new((1,), acc_results.device)
    lo = 0
    for _, yin, _ in _xy_chunks(dataset, kernel):
        from .kernels.base import to_device
        y = to_device(_as_f64(yin), torch.float64)
        c = c_mat[lo:lo + y.shape[0]]
        lo += y.shape[0]
        ops.zty_accum_c32(c, y, z_trans_y, scale, icpt)
        y_trans_y += y @ y
        ops.atb_accum_tc_c32(compressor.transform_encoded(c, scale, icpt), c, acc_results, scale, icpt)
    return y_trans_y


def _single_pass_gauss(dataset, kernel, q_mat, acc_results, ops, scratch_holder):
    """rand_nys_constructors.py:18-36: acc (M, rank) += Z^T (Z Q) -- the CG matvec with `rank` right-hand sides."""
    q_mat = q_mat.contiguous()
    tc = getattr(ops, "ztzq_accum_tc", None)
    if tc is not None and q_mat.shape[0] <= 65536 and q_mat.shape[1] >= 64:
        # wide Q: both GEMMs on the tensor cores (exact int8 digit planes, ~1e-12 of the row / column scale)
        for z in _transformed_chunks(dataset, kernel):
            tc(z, q_mat, acc_results)
        return
    for xdata, ldata in _x_chunks(dataset, kernel):
        z = kernel.transform_x(xdata, ldata)
        s = scratch_holder.get("s")
        if s is None or s.shape[0] < z.shape[0] or s.shape[1] != q_mat.shape[1]:
            s = scratch_holder["s"] = torch.empty((z.shape[0], q_mat.shape[1]), dtype=torch.float64, device=z.device)
        ops.cg_matvec_accum(z, q_mat, acc_results, s)


def _psd_eig(c_mat):
    """(eigenvalues descending, V^T rows = eigenvectors) of a symmetric PSD matrix: what svd(c_mat) returns as (s, vh)."""
    if c_mat.is_cuda:
        evals, evecs = torch.linalg.eigh((c_mat + c_mat.T) * 0.5)
        return evals.flip(0).clamp(min=0), evecs.flip(1).T.contiguous()
    _, s, vh = torch.linalg.svd(c_mat, full_matrices=False)
    return s, vh


def _tall_svd(mat):
    """(U, s) of a tall (M, rank) matrix B -- the replicated, serial step of the Nystrom build.
    On the GPU: eigh of the small Gram B^T B (s^2 and V; 6 ms at 16384 x 512 against 22 ms for cuSOLVER's Jacobi SVD and
    43-60 ms for gesvd), U = B V / s.  eigh resolves s^2 to eps * s_max^2 absolutely, so the smallest eigenvalue (which
    feeds achieved_ratio, the prefactor and the log-determinant) and the orthogonality of U carry a relative error of
    eps * s_max^2 / s_min^2: when the spectrum of s^2 spans more than 1e8 (more than ~2e-8 relative) the Jacobi SVD is
    used instead."""
    if mat.is_cuda:
        evals, evecs = torch.linalg.eigh(mat.T @ mat)
        evals, evecs = evals.flip(0), evecs.flip(1)
        if float(evals[-1]) > 1e-8 * float(evals[0]):
            s = torch.sqrt(evals)
            return (mat @ evecs) / s[None, :], s
        u, s, _ = torch.linalg.svd(mat, full_matrices=False, driver="gesvdj")
        return u, s
    u, s, _ = torch.linalg.svd(mat, full_matrices=False)
    return u, s


def initialize_srht(dataset, rank, kernel, random_state, comm, ops, compressor_cls, overlap_hook=None,
                    encoded_source=None):
    """rand_nys_constructors.py:221-296 (single pass).  `overlap_hook` is called once, after the accumulation pass and its
    all-reduces have been issued and before the serial factorisation (see ConjugateGrad.prepare_async).
    `encoded_source()` -> the shard's resident float32 encoding of Z or None (ConjugateGrad.resident_for_preconditioner): when
    the solve that follows keeps it anyway, the accumulation pass reads it instead of generating fp64 features again."""
    dev = _device_of(kernel)
    m = kernel.get_num_rffs()
    acc_results, z_trans_y = _new((rank, m), dev), _new((m,), dev)
    compressor = compressor_cls(rank, m, random_seed=random_state, device=kernel.device)
    y_trans_y = None
    c_mat = encoded_source() if encoded_source is not None else None
    if c_mat is not None:
        y_trans_y = _single_pass_srht_zty_encoded(dataset, kernel, compressor, acc_results, z_trans_y, ops, c_mat)
    if y_trans_y is None:
        acc_results.zero_()
        z_trans_y.zero_()
        y_trans_y = _single_pass_srht_zty(dataset, kernel, compressor, acc_results, z_trans_y, ops)
    comm.all_reduce_(acc_results)
    comm.all_reduce_(z_trans_y, y_trans_y)
    if overlap_hook is not None:
        overlap_hook()

    c_mat = compressor.transform_x(acc_results)
    # c_mat = Omega^T Z^T Z Omega is symmetric PSD and the reference forms Y c_mat^(-1/2) from its SVD
    # (rand_nys_constructors.py:274-283) only to take the SVD of that product: U S^2 U^T = Y c_mat^-1 Y^T.  Any factor
    # c_mat = L L^T gives the same U and S (Y L^-T differs from Y c_mat^(-1/2) by an orthogonal matrix on the right), so the
    # Cholesky factor + one triangular solve (0.6 ms at rank 512) replaces the eigendecomposition (5 ms) -- this and the tall
    # SVD below are the replicated, serial part of the build, i.e. the Amdahl term of the sharded fit.
    # A numerically singular c_mat (rank above the numerical rank of Z) takes the reference's route with its clamps.
    sym = (c_mat + c_mat.T) * 0.5
    chol, info = torch.linalg.cholesky_ex(sym)
    if int(info) == 0 and float(chol.diagonal().min()) > 1e-7 * float(chol.diagonal().max()):
        acc_results = torch.linalg.solve_triangular(chol, acc_results, upper=False).T.contiguous()
    else:
        c_s1, c_v1 = _psd_eig(c_mat)
        mask = c_s1 < 1e-14
        c_s1 = 1.0 / torch.sqrt(c_s1.clamp(min=1e-14))
        c_s1[mask] = 0
        acc_results = acc_results.T @ c_v1.T @ (c_s1[:, None] * c_v1)
    u_mat, s_mat = _tall_svd(acc_results)
    return u_mat, s_mat ** 2, z_trans_y, float(y_trans_y[0])


def initialize_srht_multipass(dataset, rank, kernel, random_state, comm, ops, compressor_cls, n_passes=2,
                              overlap_hook=None):
    """rand_nys_constructors.py:128-218 (SRHT pass, then QR + Z^T (Z Q) passes, shifted Cholesky, SVD)."""
    dev = _device_of(kernel)
    m = kernel.get_num_rffs()
    acc_results, z_trans_y = _new((rank, m), dev), _new((m,), dev)
    compressor = compressor_cls(rank, m, random_seed=random_state, device=kernel.device)
    y_trans_y = _single_pass_srht_zty(dataset, kernel, compressor, acc_results, z_trans_y, ops)
    comm.all_reduce_(acc_results)
    comm.all_reduce_(z_trans_y, y_trans_y)
    acc_results = acc_results.T.contiguous()
    q_mat, holder = None, {}
    for _ in range(n_passes - 1):
        q_mat, _ = torch.linalg.qr(acc_results)
        acc_results.zero_()
        _single_pass_gauss(dataset, kernel, q_mat, acc_results, ops, holder)
        comm.all_reduce_(acc_results)
    if overlap_hook is not None:
        overlap_hook()
    norm = float(torch.sqrt((acc_results ** 2).sum()))
    shift = float(np.spacing(norm))
    acc_results += shift * q_mat
    q_mat = torch.linalg.cholesky(q_mat.T @ acc_results)
    acc_results = torch.linalg.solve_triangular(q_mat, acc_results.T, upper=False).T
    u_mat, s_mat = _tall_svd(acc_results)
    s_mat = (s_mat ** 2 - shift).clamp(min=0)
    return u_mat, s_mat, z_trans_y, float(y_trans_y[0])


class RandNysPreconditioner:
    """rand_nys_preconditioners.py:14-72."""

    def __init__(self, kernel, dataset, max_rank, verbose=False, random_state=123, method="srht", comm=None,
                 ops=_cuda_ops, compressor_cls=None, overlap_hook=None, encoded_source=None):
        if method not in ["srht_2", "srht_3", "srht"]:
            raise RuntimeError("Unknown method supplied for tuning preconditioner construction.")
        comm = comm or RowShardComm()
        if compressor_cls is None:
            from .kernels import SRHTCompressor as compressor_cls
        if method.startswith("srht_"):
            self.u_mat, self.eig, self.z_trans_y, self.y_trans_y = initialize_srht_multipass(
                dataset, max_rank, kernel, random_state, comm, ops, compressor_cls, int(method.split("_")[1]),
                overlap_hook)
        else:
            self.u_mat, self.eig, self.z_trans_y, self.y_trans_y = initialize_srht(
                dataset, max_rank, kernel, random_state, comm, ops, compressor_cls, overlap_hook, encoded_source)
        lambda_ = kernel.get_lambda()
        min_eig = float(self.eig.min())
        self.eig = self.eig + lambda_ ** 2
        self.inv_eig = self.eig.clone()
        mask = self.inv_eig > 1e-14
        self.inv_eig[mask] = 1 / self.inv_eig[mask]
        self.inv_eig[~mask] = 0.0
        self.achieved_ratio = min_eig / lambda_ ** 2
        self.prefactor = float(min_eig + lambda_ ** 2)
        self.device = kernel.device

    def batch_matvec(self, xvec):
        xprod = self.u_mat.T @ xvec
        xprod1 = self.u_mat @ (self.inv_eig[:, None] * self.prefactor * xprod)
        xprod2 = xvec - (self.u_mat @ xprod)
        return xprod2 + xprod1

    def rev_batch_matvec(self, xvec):
        """Matvec with the NON-inverted preconditioner (rand_nys_preconditioners.py:75-81)."""
        xprod = self.u_mat.T @ xvec
        xprod1 = self.u_mat @ (self.eig[:, None] * xprod) / self.prefactor
        xprod2 = xvec - (self.u_mat @ xprod)
        return xprod2 + xprod1

    def get_zty(self):
        return self.z_trans_y

    def get_yty(self):
        return self.y_trans_y

    def get_rank(self):
        return self.inv_eig.shape[0]

    def get_logdet(self):
        """log det of the preconditioner (rand_nys_preconditioners.py:97-103)."""
        logdet = 1 + (self.eig - self.prefactor) / self.prefactor
        return float(torch.log(logdet.clamp(min=1e-12)).sum())

    def matvec_for_sampling(self, xvec):
        """P^(1/2) x: turns standard-normal probes into samples with covariance P (rand_nys_preconditioners.py:106-119)."""
        eigvals = torch.sqrt(self.eig.clamp(min=0))
        prefactor = float(np.sqrt(1 / self.prefactor))
        xprod = self.u_mat.T @ xvec
        xprod1 = self.u_mat @ (eigvals[:, None] * prefactor * xprod)
        xprod2 = xvec - (self.u_mat @ xprod)
        return xprod1 + xprod2


# ----------------------------------------------------------------------------------------- predict
def predict(kernel, weights, input_x, sequence_lengths=None, chunk_size=2000, trainy_mean=0.0, trainy_std=1.0,
            var=None, variance_rffs=None, ops=_cuda_ops):
    """xgp_regression.py:105-134: (Z * w).sum(1), de-standardised; optional exact variance (:120-126).

    Mean-only prediction with a fixed-vector SORF kernel never writes Z: every chunk goes through the fused
    featurise-dot kernel (`rbf_zv`: t = Z w with the features regenerated in registers), so inference
    moves 4 D bytes in and 8 bytes out per row instead of 8 M.  With a variance request (which needs the first
    `variance_rffs` columns of Z) or any other kernel family the chunk's Z is materialised as in the reference."""
    preds, pvar = [], []
    lambda_ = kernel.get_lambda()

    def chunks():
        for i in range(0, input_x.shape[0], chunk_size):
            cutoff = min(i + chunk_size, input_x.shape[0])
            yield input_x[i:cutoff, ...], None if sequence_lengths is None else sequence_lengths[i:cutoff]

    if var is None and sequence_lengths is None and kernel.device == "cuda" and _fused_ok(kernel, ops):
        w = weights.to(torch.float64).contiguous()
        for xc, _ in chunks():
            xs = kernel.stage_scaled(xc)
            t = torch.zeros(xs.shape[0], dtype=torch.float64, device=xs.device)
            kernel.fused_zv(xs, w, t)
            preds.append(t)
        return torch.cat(preds).cpu().numpy() * trainy_std + trainy_mean

    stream = getattr(kernel, "transform_stream", None)
    zs = stream(chunks()) if stream is not None and kernel.device == "cuda" else \
        (kernel.transform_x(xc, sl) for xc, sl in chunks())
    for z in zs:
        preds.append((z * weights[None, :]).sum(dim=1))
        if var is not None:
            zv = z[:, :variance_rffs]
            pv = (var @ zv.T).T
            pvar.append(lambda_ ** 2 + lambda_ ** 2 * (zv * pv).sum(dim=1))
    preds = torch.cat(preds).cpu().numpy() * trainy_std + trainy_mean
    if var is None:
        return preds
    pvar = torch.cat(pvar).cpu().numpy()
    pvar[pvar < 0] = 0
    return preds, pvar * trainy_std ** 2
