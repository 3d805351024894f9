"""ctypes loader for libxgpr_b200.so (the C ABI of include/xgpr_b200.h).

There is NO fallback: if the library has not been built, importing this module raises, and every
op in xgpr_b200 fails loudly (the product path never routes through oracle/ or the CPU).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("XGPR_B200_LIB") or os.path.join(_HERE, "lib", "libxgpr_b200.so")   # override: A/B of two builds

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -m xgpr_b200.build_native` "
        "(or __graft_entry__.build()). xgpr_b200 has no CPU / PyTorch fallback.")

lib = ctypes.CDLL(LIB_PATH)

_i64, _int, _dbl, _ptr = ctypes.c_int64, ctypes.c_int, ctypes.c_double, ctypes.c_void_p

# name -> argtypes (all return int unless noted); mirrors include/xgpr_b200.h one to one
SIGNATURES = {
    "xgpr_b200_fht2d": [_ptr, _i64, _i64, _ptr],
    "xgpr_b200_scale_input": [_ptr, _i64, _i64, _ptr, _dbl, _ptr],
    "xgpr_b200_srht": [_ptr, _i64, _i64, _ptr, _i64, _ptr],
    "xgpr_b200_rbf_feature_gen": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _int, _ptr],
    "xgpr_b200_rbf_zv": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _int, _ptr, _ptr, _ptr],
    "xgpr_b200_rbf_ztv": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _int, _ptr, _ptr, _ptr],
    "xgpr_b200_rbf_zzv": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _int, _ptr, _ptr, _ptr],
    "xgpr_b200_rbf_grad": [_ptr, _i64, _i64, _ptr, _ptr, _i64, _ptr, _i64, _ptr, _i64, _dbl, _int, _ptr],
    "xgpr_b200_ard_grad": [_ptr, _i64, _i64, _ptr, _ptr, _i64, _ptr, _ptr, _ptr, _i64, _int, _ptr],
    "xgpr_b200_ard_grad_overwrite": [_ptr, _i64, _i64, _ptr, _ptr, _i64, _ptr, _ptr, _ptr, _i64, _int, _ptr],
    "xgpr_b200_conv1d_maxpool": [_ptr, _i64, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _ptr, _int, _ptr],
    "xgpr_b200_conv1d_fgen": [_ptr, _i64, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _ptr, _int, _int, _ptr],
    "xgpr_b200_conv1d_grad": [_ptr, _i64, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _ptr, _ptr, _dbl,
                              _int, _int, _ptr],
}
F64_ONLY = {
    "xgpr_b200_zty_accum_f64": [_ptr, _i64, _i64, _ptr, _ptr, _ptr],
    "xgpr_b200_ztz_accum_f64": [_ptr, _i64, _i64, _ptr, _ptr],
    "xgpr_b200_ztz_accum_tc_f64": [_ptr, _i64, _i64, _ptr, _int, _int, _int, _ptr, _i64, _ptr],
    "xgpr_b200_symmetrize_upper_f64": [_ptr, _i64, _ptr],
    "xgpr_b200_atb_accum_tc_f64": [_ptr, _ptr, _i64, _i64, _i64, _ptr, _int, _int, _ptr, _i64, _ptr],
    "xgpr_b200_ztzq_accum_tc_f64": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _int, _int, _ptr, _i64, _ptr],
    "xgpr_b200_atb_accum_f64": [_ptr, _ptr, _i64, _i64, _i64, _ptr, _ptr],
    "xgpr_b200_cg_matvec_accum_f64": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _ptr, _ptr],
    "xgpr_b200_cg_matvec_accum_static_f64": [_ptr, _i64, _i64, _ptr, _i64, _ptr, _ptr],
}

EXPORTED = ["xgpr_b200_abi_version", "xgpr_b200_last_error", "xgpr_b200_launch_count"]
for _name, _args in SIGNATURES.items():
    for _sfx in ("_f32", "_f64"):
        _fn = getattr(lib, _name + _sfx)
        _fn.argtypes = _args
        _fn.restype = _int
        EXPORTED.append(_name + _sfx)
for _name, _args in F64_ONLY.items():
    _fn = getattr(lib, _name)
    _fn.argtypes = _args
    _fn.restype = _int
    EXPORTED.append(_name)

lib.xgpr_b200_zty_accum_c32.argtypes = [_ptr, _i64, _i64, _ptr, _ptr, _dbl, _int, _ptr]
lib.xgpr_b200_zty_accum_c32.restype = _int
lib.xgpr_b200_srht_sample_c32.argtypes = [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _dbl, _int, _ptr]
lib.xgpr_b200_srht_sample_c32.restype = _int
lib.xgpr_b200_atb_accum_tc_c32.argtypes = [_ptr, _ptr, _i64, _i64, _i64, _ptr, _dbl, _int, _int, _int, _ptr, _i64, _ptr]
lib.xgpr_b200_atb_accum_tc_c32.restype = _int
EXPORTED += ["xgpr_b200_zty_accum_c32", "xgpr_b200_srht_sample_c32", "xgpr_b200_atb_accum_tc_c32"]
lib.xgpr_b200_cg_matvec_accum_c32.argtypes = [_ptr, _i64, _i64, _ptr, _i64, _ptr, _ptr, _dbl, _int, _ptr]
lib.xgpr_b200_cg_matvec_accum_c32.restype = _int
EXPORTED.append("xgpr_b200_cg_matvec_accum_c32")
lib.xgpr_b200_cg_matvec_static_max_elems.argtypes = []
lib.xgpr_b200_cg_matvec_static_max_elems.restype = _i64
EXPORTED.append("xgpr_b200_cg_matvec_static_max_elems")
lib.xgpr_b200_ztz_tc_workspace_bytes.argtypes = [_i64, _i64, _int]
lib.xgpr_b200_ztz_tc_workspace_bytes.restype = _i64
EXPORTED.append("xgpr_b200_ztz_tc_workspace_bytes")
lib.xgpr_b200_atb_tc_workspace_bytes.argtypes = [_i64, _i64, _i64, _int]
lib.xgpr_b200_atb_tc_workspace_bytes.restype = _i64
EXPORTED.append("xgpr_b200_atb_tc_workspace_bytes")
lib.xgpr_b200_ztzq_tc_workspace_bytes.argtypes = [_i64, _i64, _i64, _int]
lib.xgpr_b200_ztzq_tc_workspace_bytes.restype = _i64
EXPORTED.append("xgpr_b200_ztzq_tc_workspace_bytes")
lib.xgpr_b200_rbf_feature_gen_c32.argtypes = [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _i64, _ptr]
lib.xgpr_b200_rbf_feature_gen_c32.restype = _int
EXPORTED.append("xgpr_b200_rbf_feature_gen_c32")
lib.xgpr_b200_unscale_to_f32.argtypes = [_ptr, _i64, ctypes.c_double, _ptr, _ptr]
lib.xgpr_b200_unscale_to_f32.restype = _int
EXPORTED.append("xgpr_b200_unscale_to_f32")
lib.xgpr_b200_cres_zzv_supported.argtypes = [_i64]
lib.xgpr_b200_cres_zzv_supported.restype = _int
EXPORTED.append("xgpr_b200_cres_zzv_supported")
lib.xgpr_b200_cres_zzv_f32.argtypes = [_ptr, _i64, _i64, _ptr, _ptr, ctypes.c_double, _int, _ptr]
lib.xgpr_b200_cres_zzv_f32.restype = _int
EXPORTED.append("xgpr_b200_cres_zzv_f32")
lib.xgpr_b200_cres_zzv_gated_f32.argtypes = [_ptr, _i64, _i64, _ptr, _ptr, ctypes.c_double, _int, _ptr, _ptr]
lib.xgpr_b200_cres_zzv_gated_f32.restype = _int
lib.xgpr_b200_cg_update_a_f64.argtypes = [_ptr, _ptr, _ptr, _ptr, _i64, _dbl, _ptr, _ptr, _i64, _ptr]
lib.xgpr_b200_cg_update_a_f64.restype = _int
lib.xgpr_b200_cg_update_b_f64.argtypes = [_ptr, _ptr, _ptr, _ptr, _i64, _dbl, _int, _ptr, _ptr]
lib.xgpr_b200_cg_update_b_f64.restype = _int
EXPORTED += ["xgpr_b200_cres_zzv_gated_f32", "xgpr_b200_cg_update_a_f64", "xgpr_b200_cg_update_b_f64"]
lib.xgpr_b200_srht_sample_supported.argtypes = [_i64, _int]
lib.xgpr_b200_srht_sample_supported.restype = _int
for _sfx in ("_f32", "_f64"):
    _fn = getattr(lib, "xgpr_b200_srht_sample" + _sfx)
    _fn.argtypes = [_ptr, _i64, _i64, _ptr, _i64, _ptr, _i64, _ptr, _ptr]
    _fn.restype = _int
EXPORTED += ["xgpr_b200_srht_sample_supported", "xgpr_b200_srht_sample_f32", "xgpr_b200_srht_sample_f64"]
lib.xgpr_b200_rbf_zzv_supported.argtypes = [_i64, _i64]
lib.xgpr_b200_rbf_zzv_supported.restype = _int
EXPORTED.append("xgpr_b200_rbf_zzv_supported")
lib.xgpr_b200_tc_profile.argtypes = [_int]
lib.xgpr_b200_tc_profile.restype = _int
lib.xgpr_b200_tc_last_gemm_ms.argtypes = []
lib.xgpr_b200_tc_last_gemm_ms.restype = _dbl
EXPORTED += ["xgpr_b200_tc_profile", "xgpr_b200_tc_last_gemm_ms"]
lib.xgpr_b200_set_accurate_sincos.argtypes = [_int]
lib.xgpr_b200_set_accurate_sincos.restype = _int
EXPORTED.append("xgpr_b200_set_accurate_sincos")
lib.xgpr_b200_abi_version.restype = _int
lib.xgpr_b200_last_error.restype = ctypes.c_char_p
lib.xgpr_b200_launch_count.restype = _i64


def last_error() -> str:
    return lib.xgpr_b200_last_error().decode("utf-8", "replace")


def launch_count() -> int:
    return int(lib.xgpr_b200_launch_count())


def check(rc: int) -> None:
    """Maps the C return code to the reference's exception type (std::runtime_error -> RuntimeError)."""
    if rc != 0:
        raise RuntimeError(last_error())
