"""ctypes binding of the C-ABI in include/dli.h.

The shared library is the only compute path of this package.  There is no CPU
or eager-PyTorch fallback: if `libdli_b200.so` is missing, loading fails loudly
with the build command to run.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdli_b200.so")

DLI_OK = 0
EPI_NONE, EPI_BIAS, EPI_RESIDUAL, EPI_GELU_NEW = 0, 1, 2, 4

_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t

# name -> (restype, argtypes); mirrors include/dli.h one to one
SIGNATURES = {
    "dli_last_error": (C.c_char_p, []),
    "dli_version": (_i, []),
    "dli_sm_count": (_i, []),
    "dli_embedding_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "dli_add_rmsnorm_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _f, _vp]),
    "dli_add_layernorm_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _f, _vp]),
    "dli_rope_kv_append_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "dli_paged_attn_decode_scratch_bytes": (_sz, [_i, _i, _i, _i]),
    "dli_paged_attn_decode_bf16": (
        _i,
        [_vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _vp, _sz, _i, _i, _i, _i, _i, _i, _f, _i, _vp],
    ),
    "dli_paged_attn_decode_static_max_bytes": (_sz, []),
    "dli_paged_attn_decode_static_bf16": (
        _i,
        [_vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _f, _i, _vp],
    ),
    "dli_paged_attn_decode_fused_bf16": (
        _i,
        [_vp, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _sz, _i, _i, _i, _i, _i, _i, _f, _i,
         _vp, _sz, _vp, _sz, _vp],
    ),
    "dli_attn_prefill_paged_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "dli_sample_rows_bf16": (_i, [_vp, _i, _vp, _i, _i, _f, _i, _f, _vp, _vp, _vp, _vp]),
    "dli_set_pdl": (_i, [_i]),
    "dli_gemm_splitk_scratch_bytes": (_sz, [_i, _i]),
    "dli_gemm_splitk_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _sz, _vp]),
    "dli_gemm_splitk_gated_silu_bf16": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _sz, _vp]),
    "dli_rmsnorm_gemv_bf16": (_i, [_vp, _vp, _f, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "dli_rmsnorm_gemv_gated_silu_bf16": (_i, [_vp, _vp, _f, _vp, _vp, _i, _i, _i, _vp]),
    "dli_attn_prefill_bf16": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "dli_attn_prefill_tc_bf16": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "dli_gemv_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "dli_gemm_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "dli_linear_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "dli_gemv_gated_silu_bf16": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "dli_gemv_mma_bf16": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "dli_gemv_mma_gated_silu_bf16": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "dli_gemm_gated_silu_bf16": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "dli_linear_gated_silu_bf16": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "dli_gemm_qkv_rope_append_bf16": (_i, [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i, _vp]),
    "dli_argmax_bf16": (_i, [_vp, _i, _vp, _i, _i, _vp]),
    "dli_argmax_candidate_bf16": (_i, [_vp, _i, _i, _vp, _i, _i, _vp]),
    "dli_argmax_merge": (_i, [_vp, _i, _i, _vp, _vp]),
    "dli_argmax_merge_strided": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "dli_gemv_argmax_scratch_bytes": (_sz, [_i]),
    "dli_gemv_argmax_bf16": (_i, [_vp, _vp, _f, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "dli_sample_bf16": (_i, [_vp, _i, _vp, _i, _i, _f, _i, _f, _vp, _vp, _vp, _vp, _i, _vp]),
    "dli_decode_advance": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp]),
}


class DliError(RuntimeError):
    """A C-ABI entry point returned a non-zero code."""

    def __init__(self, fn: str, code: int, message: str):
        super().__init__(f"{fn} failed ({code}): {message}")
        self.fn, self.code, self.message = fn, code, message


_lib = None


def load() -> C.CDLL:
    """Load (once) and return the CDLL with argtypes set."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the CUDA extension is the only compute path of this "
            "package (no CPU fallback). Build it with "
            "`python distributed-llm-inferencing_b200/build.py` or `python -c 'import __graft_entry__ as g; g.build()'`."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().dli_last_error().decode("utf-8", "replace")


def check(fn: str, code: int) -> None:
    if code != DLI_OK:
        raise DliError(fn, code, last_error())
