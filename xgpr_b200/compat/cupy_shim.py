"""A minimal `cupy` stand-in backed by torch CUDA tensors -- exactly the surface the reference front end touches.

SURVEY.md 8(b) enumerated it by grep over /root/reference/src/xGPR: zeros, asarray, asnumpy, float32 / float64, ndarray,
get_default_memory_pool().free_all_blocks(), linalg.{svd, cholesky, qr, pinv, norm}, log, sqrt, cos, sin, diag, eye, stack,
vstack, concatenate, transpose, ascontiguousarray, cupyx.scipy.linalg.solve_triangular; and on arrays: .T, @, .sum(axis=),
.clip(min=), .flat[::n+1] +=, .copy(), .astype, slicing / slice assignment, arithmetic with NumPy scalars.

`ndarray` wraps a torch.Tensor (`._t`).  Three expressions of the reference's L4 loops are routed to this package's own
consumers of Z instead of a library GEMM (they are the hot path):
    Z.T @ Z  (same array on both sides)   -> ext.ztz_accum_tc  (tcgen05 split-int8 Gram)      exact_nmll_calcs.py:70
    Z.T @ y  (fp64 matrix x vector)       -> ext.zty_accum                                    exact_nmll_calcs.py:69
    A.T @ Z  (fp64, wide)                 -> ext.atb_accum_tc                                 rand_nys_constructors.py:119
Everything else is torch (cuBLAS / cuSOLVER), as it was CuPy's cuBLAS / cuSOLVER in the reference.

Type promotion follows NumPy (NEP 50), which CuPy follows: Python scalars are weak, NumPy scalars and arrays are strong.
`x *= np.float64(s)` on a float32 array -- the reference's `input_x *= self.hyperparams[1]`
(sorf_kernel_baseclass.py:117) -- therefore forms the product in double and rounds once; it is routed to the
`scale_input` kernel that does exactly that.

This is a compatibility layer for running the UNMODIFIED reference front end on libxgpr_b200.so; the package's own host
layer (xgpr_b200.kernels / fitting / scoring) does not use it.
"""
import numbers

import numpy as np
import torch

float32, float64, int32, int8, int64 = np.float32, np.float64, np.int32, np.int8, np.int64

_NP2T = {np.dtype(np.float32): torch.float32, np.dtype(np.float64): torch.float64, np.dtype(np.int32): torch.int32,
         np.dtype(np.int64): torch.int64, np.dtype(np.int8): torch.int8, np.dtype(np.bool_): torch.bool,
         np.dtype(np.uint8): torch.uint8, np.dtype(np.int16): torch.int16}
_T2NP = {v: k for k, v in _NP2T.items()}


def _tdtype(dtype):
    if dtype is None:
        return None
    if isinstance(dtype, torch.dtype):
        return dtype
    return _NP2T[np.dtype(dtype)]


def _device():
    if not torch.cuda.is_available():
        raise RuntimeError("the cupy stand-in of xgpr_b200 needs a CUDA device (device='cuda' has no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device())


def _unwrap(x):
    return x._t if isinstance(x, ndarray) else x


def _key(k):
    """Index objects may contain stand-in arrays or NumPy index arrays."""
    if isinstance(k, tuple):
        return tuple(_key(i) for i in k)
    if isinstance(k, ndarray):
        return k._t
    if isinstance(k, np.ndarray):
        return torch.from_numpy(k).to(_device())
    if isinstance(k, np.integer):
        return int(k)
    return k


class _Flat:
    """`arr.flat[sl]` read / write on a C-contiguous array (the reference adds to the diagonal with
    `a.flat[::n+1] += c`, exact_fitting_toolkit.py:34)."""

    def __init__(self, t):
        if not t.is_contiguous():
            raise ValueError("flat needs a C-contiguous array")
        self._v = t.view(-1)

    def __getitem__(self, k):
        return ndarray(self._v[_key(k)])

    def __setitem__(self, k, value):
        dst = self._v[_key(k)]
        src = _unwrap(value)
        if isinstance(src, torch.Tensor) and src.data_ptr() == dst.data_ptr() and src.shape == dst.shape:
            return                                   # the in-place op already wrote through the view
        dst.copy_(src) if isinstance(src, torch.Tensor) else dst.fill_(src)


class ndarray:
    __array_priority__ = 1000.0                      # NumPy scalars / arrays defer to our reflected operators
    __hash__ = None

    def __init__(self, t):
        self._t = t

    # ---- descriptors
    @property
    def shape(self):
        return tuple(self._t.shape)

    @property
    def ndim(self):
        return self._t.dim()

    @property
    def size(self):
        return self._t.numel()

    @property
    def dtype(self):
        return _T2NP[self._t.dtype]

    @property
    def T(self):
        return ndarray(self._t.T if self._t.dim() == 2 else self._t.permute(*reversed(range(self._t.dim()))))

    @property
    def flat(self):
        return _Flat(self._t)

    @property
    def flags(self):
        return {"C_CONTIGUOUS": self._t.is_contiguous()}

    @property
    def device(self):
        return self._t.device

    @property
    def __cuda_array_interface__(self):
        return self._t.__cuda_array_interface__

    def __dlpack__(self, *args, **kwargs):
        return self._t.__dlpack__(*args, **kwargs)

    def __dlpack_device__(self):
        return self._t.__dlpack_device__()

    def __len__(self):
        return self._t.shape[0]

    def __float__(self):
        return float(self._t)

    def __int__(self):
        return int(self._t)

    def __bool__(self):
        return bool(self._t)

    def __repr__(self):
        return "xgpr_b200.compat.ndarray(" + repr(self._t) + ")"

    def __iter__(self):
        for i in range(self._t.shape[0]):
            yield ndarray(self._t[i])

    # ---- indexing
    def __getitem__(self, k):
        return ndarray(self._t[_key(k)])

    def __setitem__(self, k, value):
        v = _unwrap(value)
        if isinstance(v, np.ndarray):
            v = torch.from_numpy(v).to(self._t.device)
        self._t[_key(k)] = v

    # ---- arithmetic (NumPy promotion: Python scalars weak, NumPy scalars / arrays strong)
    def _operand(self, other):
        """-> (tensor or Python scalar, result dtype to compute in or None)."""
        if isinstance(other, ndarray):
            return other._t, None
        if isinstance(other, np.ndarray):
            return torch.from_numpy(other).to(self._t.device), None
        if isinstance(other, np.generic):
            if isinstance(other, np.floating) and self._t.dtype.is_floating_point:
                want = torch.promote_types(self._t.dtype, _NP2T[other.dtype])
                return float(other), (want if want != self._t.dtype else None)
            return other.item(), None
        if isinstance(other, (numbers.Number, torch.Tensor)):
            return other, None
        return NotImplemented, None

    def _binary(self, other, fn, reflected=False):
        o, want = self._operand(other)
        if o is NotImplemented:
            return NotImplemented
        a = self._t if want is None else self._t.to(want)
        return ndarray(fn(o, a) if reflected else fn(a, o))

    def _inplace(self, other, fn_name):
        o, want = self._operand(other)
        if o is NotImplemented:
            return NotImplemented
        if want is not None:                          # e.g. float32 array op= np.float64 scalar: double, one rounding
            if fn_name == "mul_" and self._t.is_contiguous() and self._t.is_cuda:
                from .. import ext
                ext.scale_input_(self._t, float(o))
            else:
                res = getattr(self._t.to(want), fn_name)(o)
                self._t.copy_(res)
            return self
        if isinstance(o, torch.Tensor) and o.dtype != self._t.dtype and \
                torch.promote_types(o.dtype, self._t.dtype) != self._t.dtype:
            res = getattr(self._t.to(torch.promote_types(o.dtype, self._t.dtype)), fn_name)(o)
            self._t.copy_(res)                        # same-kind cast back, as NumPy's in-place ufuncs
            return self
        getattr(self._t, fn_name)(o)
        return self

    def __add__(self, o): return self._binary(o, torch.add)
    def __radd__(self, o): return self._binary(o, torch.add, True)
    def __sub__(self, o): return self._binary(o, torch.sub)
    def __rsub__(self, o): return self._binary(o, lambda a, b: torch.sub(torch.as_tensor(a, device=b.device), b) if not isinstance(a, torch.Tensor) else torch.sub(a, b), True)
    def __mul__(self, o): return self._binary(o, torch.mul)
    def __rmul__(self, o): return self._binary(o, torch.mul, True)
    def __truediv__(self, o): return self._binary(o, torch.true_divide)
    def __rtruediv__(self, o): return self._binary(o, lambda a, b: torch.true_divide(torch.as_tensor(a, device=b.device, dtype=b.dtype if b.dtype.is_floating_point else None), b) if not isinstance(a, torch.Tensor) else torch.true_divide(a, b), True)
    def __pow__(self, o): return self._binary(o, torch.pow)
    def __neg__(self): return ndarray(-self._t)
    def __abs__(self): return ndarray(self._t.abs())
    def __iadd__(self, o): return self._inplace(o, "add_")
    def __isub__(self, o): return self._inplace(o, "sub_")
    def __imul__(self, o): return self._inplace(o, "mul_")
    def __itruediv__(self, o): return self._inplace(o, "div_")
    def __lt__(self, o): return self._binary(o, torch.lt)
    def __le__(self, o): return self._binary(o, torch.le)
    def __gt__(self, o): return self._binary(o, torch.gt)
    def __ge__(self, o): return self._binary(o, torch.ge)
    def __eq__(self, o): return self._binary(o, torch.eq)
    def __ne__(self, o): return self._binary(o, torch.ne)
    def __invert__(self): return ndarray(~self._t)

    # ---- products: the hot-path expressions go to this package's kernels
    def __matmul__(self, other):
        o = _unwrap(other)
        if isinstance(o, np.ndarray):
            o = torch.from_numpy(o).to(self._t.device)
        a = self._t
        routed = _routed_matmul(a, o)
        return ndarray(routed if routed is not None else torch.matmul(a, o))

    def __rmatmul__(self, other):
        o = other
        if isinstance(o, np.ndarray):
            o = torch.from_numpy(o).to(self._t.device)
        return ndarray(torch.matmul(o, self._t))

    # ---- methods
    def sum(self, axis=None, keepdims=False):
        return ndarray(self._t.sum() if axis is None else self._t.sum(dim=axis, keepdim=keepdims))

    def mean(self, axis=None):
        return ndarray(self._t.mean() if axis is None else self._t.mean(dim=axis))

    def max(self, axis=None):
        return ndarray(self._t.max() if axis is None else self._t.max(dim=axis).values)

    def min(self, axis=None):
        return ndarray(self._t.min() if axis is None else self._t.min(dim=axis).values)

    def copy(self):
        return ndarray(self._t.clone(memory_format=torch.contiguous_format))

    def astype(self, dtype, copy=True):
        td = _tdtype(dtype)
        if not copy and td == self._t.dtype:
            return self
        return ndarray(self._t.to(td, copy=True))

    def reshape(self, *shape):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list)):
            shape = tuple(shape[0])
        return ndarray(self._t.reshape(*shape))

    def clip(self, min=None, max=None):
        return ndarray(self._t.clamp(min=min, max=max))

    def get(self):
        return self._t.detach().cpu().numpy()

    def item(self):
        return self._t.item()

    def tolist(self):
        return self._t.tolist()

    def fill(self, v):
        self._t.fill_(v)

    def flatten(self):
        return ndarray(self._t.flatten())

    def transpose(self, *axes):
        if not axes:
            return self.T
        if len(axes) == 1 and isinstance(axes[0], (tuple, list)):
            axes = tuple(axes[0])
        return ndarray(self._t.permute(*axes))


def _routed_matmul(a, o):
    """A.T @ B with A, B row-major fp64 CUDA matrices -> this package's consumers of Z (None: not a routed shape)."""
    if not (isinstance(o, torch.Tensor) and a.is_cuda and a.dtype == torch.float64 and o.dtype == torch.float64
            and a.dim() == 2 and a.T.is_contiguous() and a.shape[1] >= 1):
        return None
    from .. import ext
    base = a.T                                       # (n, m) row-major: the chunk's features
    n, m = base.shape
    if o.dim() == 1 and o.is_contiguous() and o.shape[0] == n:
        out = torch.zeros(m, dtype=torch.float64, device=a.device)
        ext.zty_accum(base, o, out)
        return out
    if o.dim() == 2 and o.is_contiguous() and o.shape[0] == n:
        if o.data_ptr() == base.data_ptr() and o.shape == base.shape and m >= 256 and m % 8 == 0:
            out = torch.zeros((m, m), dtype=torch.float64, device=a.device)
            ext.ztz_accum_tc(base, out)                                   # upper triangle + mirror
            return out
        if o.shape[1] >= 256 and o.shape[1] % 8 == 0 and m >= 16:
            out = torch.zeros((m, o.shape[1]), dtype=torch.float64, device=a.device)
            ext.atb_accum_tc(base, o, out)
            return out
    return None


# ------------------------------------------------------------------------------------------ module-level functions
def zeros(shape, dtype=float64):
    return ndarray(torch.zeros(shape, dtype=_tdtype(dtype), device=_device()))


def empty(shape, dtype=float64):
    return ndarray(torch.empty(shape, dtype=_tdtype(dtype), device=_device()))


def ones(shape, dtype=float64):
    return ndarray(torch.ones(shape, dtype=_tdtype(dtype), device=_device()))


def asarray(a, dtype=None):
    if isinstance(a, ndarray):
        return a if dtype is None or _tdtype(dtype) == a._t.dtype else a.astype(dtype)
    if isinstance(a, torch.Tensor):
        return ndarray(a.to(_device(), dtype=_tdtype(dtype)))
    arr = np.ascontiguousarray(a) if dtype is None else np.ascontiguousarray(a, dtype=dtype)
    return ndarray(torch.from_numpy(arr).to(_device()))


array = asarray


def asnumpy(a):
    if isinstance(a, ndarray):
        return a.get()
    return np.asarray(a)


def ascontiguousarray(a, dtype=None):
    t = _unwrap(asarray(a))
    if dtype is not None:
        t = t.to(_tdtype(dtype))
    return ndarray(t.contiguous())


def _unary(fn):
    def f(a):
        return ndarray(fn(_unwrap(asarray(a))))
    return f


log, sqrt, cos, sin, exp, abs = (_unary(f) for f in (torch.log, torch.sqrt, torch.cos, torch.sin, torch.exp, torch.abs))


def diag(a):
    return ndarray(torch.diag(_unwrap(a)))


def eye(n, dtype=float64):
    return ndarray(torch.eye(n, dtype=_tdtype(dtype), device=_device()))


def stack(arrays, axis=0):
    return ndarray(torch.stack([_unwrap(a) for a in arrays], dim=axis))


def vstack(arrays):
    return ndarray(torch.vstack([_unwrap(a) for a in arrays]))


def concatenate(arrays, axis=0):
    return ndarray(torch.cat([_unwrap(a) for a in arrays], dim=axis))


def transpose(a, axes=None):
    return a.transpose() if axes is None else a.transpose(axes)


def sum(a, axis=None):                               # noqa: A001  (mirrors cupy.sum)
    return a.sum(axis=axis)


class _MemoryPool:
    def free_all_blocks(self):
        """CuPy's pool is torch's caching allocator here; the reference calls this after every large pass."""
        return None

    def used_bytes(self):
        return torch.cuda.memory_allocated()


_POOL = _MemoryPool()


def get_default_memory_pool():
    return _POOL


class _Linalg:
    @staticmethod
    def svd(a, full_matrices=True):
        u, s, vh = torch.linalg.svd(_unwrap(a), full_matrices=full_matrices)
        return ndarray(u), ndarray(s), ndarray(vh)

    @staticmethod
    def cholesky(a):
        return ndarray(torch.linalg.cholesky(_unwrap(a)))

    @staticmethod
    def qr(a, mode="reduced"):
        q, r = torch.linalg.qr(_unwrap(a), mode=mode)
        return ndarray(q), ndarray(r)

    @staticmethod
    def pinv(a):
        return ndarray(torch.linalg.pinv(_unwrap(a)))

    @staticmethod
    def norm(a, axis=None):
        t = _unwrap(a)
        return ndarray(torch.linalg.norm(t) if axis is None else torch.linalg.norm(t, dim=axis))


linalg = _Linalg()


def solve_triangular(a, b, lower=False, trans=0):
    """cupyx.scipy.linalg.solve_triangular: b may be a vector or a matrix."""
    at, bt = _unwrap(a), _unwrap(b)
    if trans in (1, "T"):
        at, lower = at.T, not lower
    vec = bt.dim() == 1
    x = torch.linalg.solve_triangular(at, bt[:, None] if vec else bt, upper=not lower)
    return ndarray(x[:, 0] if vec else x)
