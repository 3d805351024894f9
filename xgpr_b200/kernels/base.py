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

    def transform_stream(self, chunks):
        """transform_x over an iterable of (input_x, sequence_length) chunks, double-buffered: the host->device copy of
        chunk i+1 is issued on a side stream before chunk i's kernels are launched, so the copy (pinned memory: fully
        asynchronous; pageable: staged by the driver) runs under the previous chunk's feature generation.  Yields exactly
        what transform_x(input_x, sequence_length) returns, in order.  This is how the L4 chunk loops of the reference
        (`for xdata, ldata in dataset.get_chunked_x_data(): kernel.transform_x(xdata, ldata)`, e.g.
        fitting_toolkit/cg_tools.py:40-44) are meant to be driven when the dataset lives on the host."""
        if self.device != "cuda":
            raise RuntimeError("xgpr_b200 generates features on device='cuda' only (no CPU fallback).")
        side = getattr(self, "_copy_stream", None)
        if side is None:
            side = self._copy_stream = torch.cuda.Stream()

        def stage(item):
            input_x, seqlen = item
            main = torch.cuda.current_stream()
            if isinstance(input_x, torch.Tensor) and input_x.is_cuda:
                return self._stage_input(input_x), seqlen, None
            with torch.cuda.stream(side):
                xin = self._stage_input(input_x)
                done = torch.cuda.Event()
                done.record(side)
            xin.record_stream(main)
            return xin, seqlen, done

        it = iter(chunks)
        try:
            nxt = stage(next(it))
        except StopIteration:
            return
        while nxt is not None:
            xin, seqlen, done = nxt
            try:
                nxt = stage(next(it))
            except StopIteration:
                nxt = None
            if done is not None:
                torch.cuda.current_stream().wait_event(done)
            xtrans = self.kernel_specific_transform(xin, self._stage_seqlen(seqlen))
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
