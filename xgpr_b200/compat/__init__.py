"""Drop-in layer for the UNMODIFIED reference front end (boundary B2).

    import xgpr_b200.compat as compat
    compat.install()                      # before `import xGPR`
    from xGPR import xGPRegression
    model = xGPRegression(..., device="cuda")      # fit / predict / exact_nmll run on libxgpr_b200.so

install() registers, in sys.modules,
  * `xGPR.xgpr_cuda_rfgen_cpp_ext` -> xgpr_b200.compat.xgpr_cuda_rfgen_cpp_ext (the eight callables of the reference's CUDA
    extension, gpu_rf_gen/xgpr_cuda_rfgen_cpp_ext.cpp:20-93, on the C ABI of include/xgpr_b200.h);
  * `cupy`, `cupyx`, `cupyx.scipy`, `cupyx.scipy.linalg` -> the torch-backed stand-in of cupy_shim.py (only when a real CuPy is
    not importable: the reference's `try: import cupy` then succeeds and its device="cuda" branches run on torch tensors).
Nothing in the reference package is patched or monkey-patched; it is imported as it lies.
"""
import importlib
import sys
import types


def install(force=False):
    from . import cupy_shim, xgpr_cuda_rfgen_cpp_ext
    sys.modules["xGPR.xgpr_cuda_rfgen_cpp_ext"] = xgpr_cuda_rfgen_cpp_ext
    have_cupy = False
    if not force:
        try:
            have_cupy = importlib.util.find_spec("cupy") is not None and "xgpr_b200" not in getattr(
                sys.modules.get("cupy"), "__name__", "")
        except (ImportError, ValueError):
            have_cupy = False
    if have_cupy and not force:
        return "xGPR.xgpr_cuda_rfgen_cpp_ext"
    sys.modules["cupy"] = cupy_shim
    cupyx = types.ModuleType("cupyx")
    scipy = types.ModuleType("cupyx.scipy")
    linalg = types.ModuleType("cupyx.scipy.linalg")
    linalg.solve_triangular = cupy_shim.solve_triangular
    scipy.linalg = linalg
    cupyx.scipy = scipy
    sys.modules["cupyx"], sys.modules["cupyx.scipy"], sys.modules["cupyx.scipy.linalg"] = cupyx, scipy, linalg
    return "xGPR.xgpr_cuda_rfgen_cpp_ext + cupy stand-in"
