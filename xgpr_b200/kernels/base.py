"""KernelBaseclass -- the kernel-plugin interface on torch CUDA tensors.

Mirrors src/xGPR/kernels/kernel_baseclass.py (constructor :51-99, abstract hooks :103-123,
hyperparameter / bounds API :126-265, transform_x :269-299, transform_x_y :303-324,
gradient_x :328-360, gradient_x_y :364-378, device property :430-457): same method names,
argument meaning and error behaviour.  Differences, all forced by "no CuPy, no CPU fallback":

  * device="cuda" arrays are torch tensors on the current CUDA device;
  * device="cpu" only HOLDS the host-drawn parameters (NumPy) -- generating features on the CPU is
    the reference's job (the oracle); this package raises instead of falling back;
  * inputs may also be (pinned) host torch tensors or CUDA tensors, so datasets can stay resident.
"""
import abc
import functools
import inspect

import numpy as np
import torch


def _cuda_device():
    if not torch.cuda.is_available():
        raise RuntimeError("xgpr_b200 kernels need a CUDA device: there is no CPU fallback "
                           "(use the reference xGPR for device='cpu').")
    return torch.device("cuda", torch.cuda.current_device())


def to_device(arr, dtype=None):
    """NumPy array / torch tensor -> torch tensor on the current CUDA device (async from pinned memory)."""
    dev = _cuda_device()
    if isinstance(arr, np.ndarray):
        arr = torch.from_numpy(np.ascontiguousarray(arr))
    return arr.to(device=dev, dtype=dtype, non_blocking=True)


def to_host(arr):
    if isinstance(arr, torch.Tensor):
        return arr.detach().cpu().numpy()
    return arr


class KernelBaseclass(abc.ABC):
    """Base of every kernel plugin; see module docstring for the reference lines mirrored."""

    def __init__(self, num_rffs, xdim, sine_cosine_kernel=False, double_precision=False,
                 kernel_spec_parms={}):
        self.double_precision = double_precision
        if num_rffs < 2:
            raise RuntimeError("num_rffs should always be >= 2.")
        if sine_cosine_kernel:
            if not (num_rffs / 2).is_integer():
                raise RuntimeError("For sine-cosine kernels (e.g. matern, rbf) the number of random "
                                   "fourier features must be an integer multiple of two.")
            self.num_freqs = int(num_rffs / 2)
        else:
            self.num_freqs = num_rffs
        self.num_rffs = num_rffs
        self.fit_intercept = kernel_spec_parms.get("intercept", True) is not False
        self._xdim = xdim
        self.hyperparams = None
        self.bounds = None

    def __init_subclass__(cls, **kwargs):
        """Records the arguments of the OUTERMOST constructor call (the registered class's signature,
        model_baseclass.py:302-305) so that `with_num_rffs` can rebuild the kernel the way the reference's
        `num_rffs` setter does."""
        super().__init_subclass__(**kwargs)
        init = cls.__dict__.get("__init__")
        if init is None:
            return

        @functools.wraps(init)
        def wrapped(self, *args, **kw):
            if not hasattr(self, "_ctor_call"):
                self._ctor_call = (cls, args, dict(kw))
            init(self, *args, **kw)
        cls.__init__ = wrapped

    def with_num_rffs(self, num_rffs):
        """A fresh kernel of the same class, xdim, seed, precision, settings, device, hyperparameters and bounds with a
        different number of random features -- what `model.num_rffs = value` does in the reference
        (model_baseclass.py:528-540 -> _initialize_kernel :263-321)."""
        cls, args, kw = self._ctor_call
        bound = inspect.signature(cls.__init__).bind(self, *args, **kw)
        bound.arguments["num_rffs"] = int(num_rffs)
        bound.arguments["device"] = self.device
        params = dict(bound.arguments)
        params.pop("self", None)
        new = cls(**params)
        if self.bounds is not None:
            new.bounds = np.array(self.bounds, copy=True)
        if self.hyperparams is not None:
            new.set_hyperparams(np.array(self.hyperparams, copy=True), logspace=False)
        return new

    # ---- hooks every kernel implements
    @abc.abstractmethod
    def kernel_specific_transform(self, input_x, sequence_length):
        """Generate random features for device-resident input."""

    @abc.abstractmethod
    def kernel_specific_gradient(self, input_x, sequence_length):
        """Random features and d(features)/d(kernel hyperparameters)."""

    @abc.abstractmethod
    def kernel_specific_set_device(self, new_device):
        """Move kernel parameters between host (NumPy) and device (torch)."""

    @abc.abstractmethod
    def kernel_specific_set_hyperparams(self):
        """Refresh anything derived from the hyperparameters."""

    # ---- bounds / hyperparameters
    def check_bounds(self, bounds):
        if bounds.shape != self.bounds.shape:
            raise RuntimeError("The bounds supplied do not match the shape [n_hyperparams, 2] "
                               "expected by the kernel you have chosen.")

    def get_bounds(self, logspace=True):
        return np.log(self.bounds) if logspace else self.bounds

    def set_bounds(self, bounds, logspace=True):
        self.check_bounds(bounds)
        self.bounds = np.exp(bounds) if logspace else bounds

    def validate_new_datapoints(self, input_x):
        if len(input_x.shape) != len(self._xdim):
            return False
        if len(self._xdim) == 3:
            return input_x.shape[2] == self._xdim[2] and input_x.shape[1] >= 1
        return input_x.shape[1] == self._xdim[1]

    def get_hyperparams(self, logspace=True):
        return np.log(self.hyperparams) if logspace else self.hyperparams

    def set_hyperparams(self, hyperparams, logspace=True):
        self.hyperparams = np.exp(hyperparams) if logspace else hyperparams
        self.kernel_specific_set_hyperparams()

    def check_hyperparams(self, hyperparams):
        if not isinstance(hyperparams, np.ndarray):
            raise RuntimeError("The starting hyperparameters must be a numpy array.")
        if hyperparams.shape != self.hyperparams.shape:
            raise RuntimeError(f"The kernel you selected uses {self.hyperparams.shape[0]} hyperparameters. "
                               "A hyperparameter array of the incorrect shape was passed.")

    def get_lambda(self):
        return self.hyperparams[0]

    def get_num_rffs(self):
        return self.num_rffs

    def get_xdim(self):
        return self._xdim

    # ---- feature generation entry points
    def _stage_input(self, input_x):
        """A private, C-contiguous device copy in the FHT dtype (the kernels scale it in place).
        kernel_baseclass.py:277-288 -- the copy is made on the device instead of the host."""
        if self.device != "cuda":
            raise RuntimeError("xgpr_b200 generates features on device='cuda' only (no CPU fallback).")
        dtype = torch.float64 if self.double_precision else torch.float32
        dev = _cuda_device()
        if isinstance(input_x, np.ndarray):
            np_dt = np.float64 if self.double_precision else np.float32
            host = np.ascontiguousarray(input_x, dtype=np_dt)
            return torch.from_numpy(host).to(dev, non_blocking=False)
        if input_x.is_cuda:
            return input_x.to(dtype=dtype, copy=True).contiguous()
        return input_x.contiguous().to(device=dev, dtype=dtype, non_blocking=True)

    @staticmethod
    def _stage_seqlen(sequence_length):
        """seqlen stays on the HOST as int32, as in the reference (kernel_baseclass.py:290-292)."""
        if sequence_length is None:
            return None
        if isinstance(sequence_length, torch.Tensor):
            sequence_length = sequence_length.cpu().numpy()
        return np.ascontiguousarray(sequence_length.astype(np.int32, copy=False))

    def transform_x(self, input_x, sequence_length=None):
        xin = self._stage_input(input_x)
        xtrans = self.kernel_specific_transform(xin, self._stage_seqlen(sequence_length))
        if self.fit_intercept:
            xtrans[:, 0] = 1.0
        return xtrans

    def stage_stream(self, chunks, extra=None):
        """Double-buffered staging of an iterable of (input_x, sequence_length[, ...]) chunks: the host->device copy of
        chunk i+1 is issued on a side stream before chunk i is handed to the caller, so the copy (pinned memory: fully
        asynchronous; pageable: staged by the driver) runs under the caller's kernels on chunk i.  Yields
        (device copy of input_x in the FHT dtype, host int32 sequence lengths or None, extra(item) staged the same way or
        None); the current stream has already been made to wait for the copies.  `extra` maps a chunk tuple to further
        host arrays to stage alongside (e.g. y), returned as a tuple of device tensors."""
        if self.device != "cuda":
            raise RuntimeError("xgpr_b200 generates features on device='cuda' only (no CPU fallback).")
        side = getattr(self, "_copy_stream", None)
        if side is None:
            side = self._copy_stream = torch.cuda.Stream()

        def stage(item):
            input_x, seqlen = item[0], item[1]
            main = torch.cuda.current_stream()
            others = extra(item) if extra is not None else None
            on_device = isinstance(input_x, torch.Tensor) and input_x.is_cuda
            if on_device and others is None:
                return self._stage_input(input_x), seqlen, None, None
            if on_device:
                side.wait_stream(main)                   # a device-resident input may have just been produced on `main`
            with torch.cuda.stream(side):
                xin = self._stage_input(input_x)
                staged = None
                if others is not None:
                    staged = tuple(to_device(o) for o in others)
                done = torch.cuda.Event()
                done.record(side)
            xin.record_stream(main)
            for t in staged or ():
                t.record_stream(main)
            return xin, seqlen, staged, done

        it = iter(chunks)
        try:
            nxt = stage(next(it))
        except StopIteration:
            return
        while nxt is not None:
            xin, seqlen, staged, done = nxt
            try:
                nxt = stage(next(it))
            except StopIteration:
                nxt = None
            if done is not None:
                torch.cuda.current_stream().wait_event(done)
            yield xin, self._stage_seqlen(seqlen), staged

    def transform_stream(self, chunks):
        """transform_x over an iterable of (input_x, sequence_length) chunks through `stage_stream` (next chunk's H2D under
        this chunk's feature generation).  Yields exactly what transform_x(input_x, sequence_length) returns, in order.
        This is how the L4 chunk loops of the reference (`for xdata, ldata in dataset.get_chunked_x_data():
        kernel.transform_x(xdata, ldata)`, e.g. fitting_toolkit/cg_tools.py:40-44) are meant to be driven when the dataset
        lives on the host."""
        for xin, seqlen, _ in self.stage_stream(chunks):
            xtrans = self.kernel_specific_transform(xin, seqlen)
            if self.fit_intercept:
                xtrans[:, 0] = 1.0
            yield xtrans
            del xtrans                                   # one output chunk alive besides the consumer's

    def transform_x_y(self, input_x, input_y, sequence_length=None, classification=False):
        xtrans = self.transform_x(input_x, sequence_length)
        ydt = torch.int32 if classification else torch.float64
        if isinstance(input_y, np.ndarray):
            input_y = input_y.astype(np.int32 if classification else np.float64)
        return xtrans, to_device(input_y, ydt)

    def gradient_x(self, input_x, sequence_length=None):
        xin = self._stage_input(input_x)
        xtrans, xgrad = self.kernel_specific_gradient(xin, self._stage_seqlen(sequence_length))
        if self.fit_intercept:
            xtrans[:, 0] = 1.0
            if xgrad.shape[2] > 0:
                xgrad[:, 0, :] = 0.0
        return xtrans, xgrad

    def gradient_x_y(self, input_x, input_y, sequence_length=None):
        y_out = to_device(input_y)
        xtrans, dz_dsigma = self.gradient_x(input_x, sequence_length)
        return xtrans, dz_dsigma, y_out

    # ---- device switching
    @property
    def device(self):
        return self.device_

    @device.setter
    def device(self, value):
        if value not in ("cpu", "cuda"):
            raise RuntimeError("Unrecognized device supplied. Must be one of 'cpu', 'cuda'.")
        self.device_ = value
        self.kernel_specific_set_device(value)

    def _move(self, names, new_device):
        for nm in names:
            arr = getattr(self, nm, None)
            if arr is None:
                continue
            setattr(self, nm, to_device(arr) if new_device == "cuda" else to_host(arr))
