"""ctypes binding of libmodelopt_b200.so (the C ABI declared in include/modelopt_b200.h).

There is NO fallback: if the library is missing or a CUDA call fails, these raise.
"""

from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MODELOPT_B200_LIB") or os.path.join(_HERE, "lib", "libmodelopt_b200.so")

MO_F32, MO_F16, MO_BF16 = 0, 1, 2
MO_NUMERICS_GPU, MO_NUMERICS_CPU = 0, 1
MO_DIV_TRITON, MO_DIV_IEEE = MO_NUMERICS_GPU, MO_NUMERICS_CPU
MO_FLAG_NEGATIVE_IN_UNSIGNED, MO_FLAG_NEGATIVE_AMAX = 1, 2
# element / scale formats (the reference's `Types` enum order, gemm/tensor_quant_mx.h:39)
MO_FMT = {"E4M3": 0, "E5M2": 1, "INT8": 2, "E0M3": 3, "E1M2": 4, "E3M0": 5, "E2M1": 6, "E3M2": 7, "E2M3": 8, "E8M0": 9}
MO_ERR_ALIGN = -5
ABI_VERSION = 4

_p, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float

# name -> argtypes ; every function returns int except mo_last_error
SIGNATURES = {
    "mo_abi_version": [],
    "mo_device_info": [C.c_char_p, _i, C.POINTER(_i), C.POINTER(_i)],
    "mo_qdq_int": [_p, _p, _i64, _i, _p, _i64, _i64, _i, _i, _i, _p, _p],
    "mo_qdq_fp8": [_p, _p, _i64, _i, _p, _i64, _i64, _i, _p],
    "mo_qdq_nvfp4_dynamic": [_p, _p, _i64, _i64, _i, _p, _i, _i, _p],
    "mo_qdq_nvfp4_static": [_p, _p, _i64, _i, _i, _i, _p, _i, _p],
    "mo_qdq_nvfp4_static_amax": [_p, _p, _i64, _i, _i, _p, _p, _i, _f, _i, _p],
    "mo_fp4_scales": [_p, _p, _i64, _p, _i, _f, _i, _p, _p],
    "mo_fp4_cast": [_p, _p, _i64, _i, _i, _p],
    "mo_amax": [_p, _i64, _i64, _i64, _i, _p, _p],
    "mo_hist": [_p, _i64, _i, _p, _i64, _p, _p, _i, _p],
    "mo_hist_collect_begin": [_p, _i64, _i, _p, _p, _i64, _i64, _i, _i, _p],
    "mo_hist_collect_end": [_p, _i64, _i, _p, _p, _i64, _i64, _i, _i, _i, _i, _p],
    "mo_hist_collect": [_p, _i64, _i, _p, _p, _i64, _i64, _i, _i, _i, _i, _p],
    "mo_hist_percentile": [_p, _i64, C.c_double, _p, _p],
    "mo_hist_mse": [_p, _p, _i64, _i, _i, _i, _i, _i, _i, _p, _p, _p],
    "mo_hist_entropy_scratch_bytes": [_i64, _i, _i],
    "mo_hist_entropy": [_p, _i64, _i, _i, _i, _i, _p, _p, _p],
    "mo_hist_entropy_static_max_bins": [],
    "mo_hist_entropy_static": [_p, _i64, _i, _i, _i, _i, _p, _p],
    "mo_nvfp4_pack": [_p, _i64, _i, _i, _p, _p, _p, _p, _p],
    "mo_nvfp4_unpack": [_p, _p, _p, _i64, _i, _i, _i, _p, _p],
    "mo_nvfp4_fp8_scale_sweep": [_p, _i64, _i, _i, _p, _p, _i, _p],
    "mo_mse_sweep": [_p, _i64, _i64, _i64, _i, _p, _i, _i, _i, _i, _i, _i, _p, _p],
    "mo_qdq_int_block_dynamic": [_p, _p, _i64, _i, _i, _p, _i64, _i, _i, _i, _p, _p],
    "mo_qdq_int_pqs": [_p, _p, _i64, _i, _p, _i64, _i64, _i, _i, _i, _p, _p, _i64, _p],
    "mo_qdq_fp8_pqs": [_p, _p, _i64, _i, _p, _i64, _i64, _i, _p, _i64, _p],
    "mo_scale_cols_amax": [_p, _p, _i64, _i64, _i, _p, _p, _i, _p],
    "mo_awq_weight_scale": [_p, _i64, _i64, _i, _i, _p, _i, _p],
    "mo_abs_col_sum": [_p, _i64, _i64, _i, _p, _i, _p],
    "mo_sq_diff_mean": [_p, _p, _i64, _i, _p, _p],
    "mo_awq_clip_loss": [_p, _p, _i64, _i64, _i, _i64, _i64, _i, _p, _i, C.POINTER(_f), _i, _i, _i, _i, _p, _p],
    "mo_linear": [_p, _p, _p, _p, _i64, _i64, _i64, _i, _p, _p, _p],
    "mo_linear_pair": [_p, _p, _p, _p, _i64, _i64, _i64, _i, _p, _p, _p],
    "mo_linear_qdq": [_p, _p, _p, _p, _i64, _i64, _i64, _i, _p, _p, _i, _i, _i, _i, _p],
    "mo_linear_nvfp4": [_p, _p, _p, _p, _p, _p, _p, _p, _i64, _i64, _i64, _i, _p],
    "mo_linear_nvfp4_pair": [_p, _p, _p, _p, _p, _p, _p, _p, _i64, _i64, _i64, _i, _p],
    "mo_nvfp4_pack_gemm": [_p, _i64, _i64, _i, _p, _p, _p, _p],
    "mo_nvfp4_sf_swizzle_bytes": [_i64, _i64, _i],
    "mo_nvfp4_sf_swizzle": [_p, _i64, _i64, _i, _p, _p],
    "mo_qdq_mx": [_p, _p, _i64, _i64, _i, _i, _i, _i, _p, _i, _p],
    "mo_mxfp8_pack": [_p, _p, _p, _i64, _i64, _i, _i, _i, _p],
    "mo_mxfp4_pack": [_p, _p, _p, _i64, _i, _i, _i, _p],
    "mo_mx_unpack": [_p, _p, _p, _i64, _i64, _i, _i, _i, _p],
    "mo_int4_pack": [_p, _i64, _i, _i, _p, _p, _i, _p],
    "mo_int4_unpack": [_p, _p, _i64, _i, _i, _p, _p],
    "mo_nf4_pack": [_p, _i64, _i, _i, _p, _p, _i, _p],
    "mo_nf4_unpack": [_p, _p, _i64, _i, _i, _i, _p, _p],
    "mo_q8_pack": [_p, _p, _p, C.POINTER(_i64), _i, _i, _p],
    "mo_q8_unpack": [_p, _p, _p, C.POINTER(_i64), _i, _i, _i, _p],
    "mo_block_hessian_scratch_bytes": [_i64, _i64],
    "mo_block_hessian": [_p, _i64, _i64, _i, _i, _p, _p, _p],
    "mo_nvfp4_fp8_scale_sweep_hessian": [_p, _i64, _i64, _i, _i, _p, _p, _p, _i, _p],
    "mo_gptq_block_nvfp4": [_p, _p, _p, _p, _f, _p, _i64, _i, _i, _i, _i, _i, _p],
    "mo_linear_fp8": [_p, _p, _p, _p, _p, _p, _i64, _i64, _i64, _i, _p],
    "mo_linear_fp8_pair": [_p, _p, _p, _p, _p, _p, _i64, _i64, _i64, _i, _p],
    "mo_selftest_e2m1": [C.POINTER(C.c_uint64)],
    "mo_selftest_exact_div": [_i, _i, C.c_uint64, C.POINTER(C.c_uint64)],
    "mo_qdq_fp8_host": [_p, _p, _i64, _i, _f, _i],
    "mo_qdq_int_host": [_p, _p, _i64, _i, _f, _i, _i, _i],
    "mo_qdq_nvfp4_dynamic_host": [_p, _p, _i64, _i64, _i, _f, _i, _i],
    "mo_amax_host": [_p, _i64, _i, C.POINTER(_f)],
    "mo_calib_qdq_host_batch": [_i, C.POINTER(_p), C.POINTER(_p), C.POINTER(_i64), C.POINTER(_i64), _i, C.POINTER(_f), _i,
                                _i, _i],
}

_lib = None


class ModeloptB200Error(RuntimeError):
    """A C-ABI call returned a non-zero status."""


def load() -> C.CDLL:
    """Load the shared library (once).  Raises ImportError with build instructions if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python __graft_entry__.py build` "
            "(nvcc, sm_100a).  modelopt_b200 has no CPU or eager fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.argtypes = argtypes
        fn.restype = _i64 if name.endswith(("_bytes", "_max_bins")) else _i
    lib.mo_last_error.argtypes = []
    lib.mo_last_error.restype = C.c_char_p
    if lib.mo_abi_version() != ABI_VERSION:
        raise ImportError(f"ABI mismatch: library {lib.mo_abi_version()}, binding {ABI_VERSION}")
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().mo_last_error().decode(errors="replace")
        raise ModeloptB200Error(f"{what} failed (status {rc}): {msg}")
