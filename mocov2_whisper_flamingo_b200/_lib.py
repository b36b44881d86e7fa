"""ctypes binding of libgxa_b200.so (C ABI declared in include/gxa.h).

There is no CPU or PyTorch fallback: if the shared object is missing this module raises at
first use, and every entry point raises ``RuntimeError`` on a non-zero return code.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GXA_LIB: another in-tree build of the same sources (A/B variants made with GXA_BUILD_VARIANT, see build.py)
LIB_PATH = os.environ.get("GXA_LIB") or os.path.join(_HERE, "libgxa_b200.so")

GXA_PREC_BF16, GXA_PREC_FP32 = 0, 1
EPI_BF16, EPI_F32, EPI_GATE_RES, EPI_BIAS_GELU, EPI_GELU_BWD = 0, 1, 2, 3, 4

_ERR = {-1: "GXA_ERR_SHAPE", -2: "GXA_ERR_ALIGN", -3: "GXA_ERR_UNSUPPORTED", -4: "GXA_ERR_CUDA",
        -5: "GXA_ERR_DRIVER", -6: "GXA_ERR_TMAP", -7: "GXA_ERR_ARG"}

PARAM_FIELDS = ("attn_gate", "ff_gate", "in_proj_w", "in_proj_b", "out_proj_w", "out_proj_b", "attn_ln_w",
                "attn_ln_b", "ff_ln_w", "ff_ln_b", "ff0_w", "ff0_b", "ff2_w", "ff2_b")


class Dims(C.Structure):      # gxa_dims
    _fields_ = [(n, C.c_int32) for n in ("B", "Lq", "Lk", "D", "H", "precision", "has_xa", "reserved")] + \
               [("p_drop", C.c_float), ("reserved2", C.c_uint32), ("seed", C.c_uint64)]


class BlockPtrs(C.Structure):   # gxa_block_params and gxa_block_grads share this shape
    _fields_ = [(n, C.c_void_p) for n in PARAM_FIELDS]


# every symbol include/gxa.h declares: name -> (restype, argtypes)
_vp, _i, _sz = C.c_void_p, C.c_int, C.c_size_t
_dp = C.POINTER(Dims)
_bp = C.POINTER(BlockPtrs)
SIGNATURES = {
    "gxa_query_caps": (_i, [C.POINTER(_i)] * 3),
    "gxa_version": (C.c_char_p, []),
    "gxa_last_cuda_error": (C.c_char_p, []),
    "gxa_launch_count": (C.c_longlong, []),
    "gxa_dropout_multipliers_host": (_i, [C.c_float, C.c_uint64, _i, _i, _i, _vp]),
    "gxa_set_sm_limit": (C.c_int, [C.c_int]),
    "gxa_block_wcache_bytes": (_sz, [_dp]),
    "gxa_block_saved_bytes": (_sz, [_dp]),
    "gxa_block_workspace_bytes": (_sz, [_dp]),
    "gxa_xa_operand_bytes": (_sz, [_dp]),
    "gxa_block_prepare_weights": (_i, [_dp, _bp, _vp, _vp]),
    "gxa_mask_prepare": (_i, [_vp, _vp, _i, _i, _vp]),
    "gxa_xa_prepare": (_i, [_dp, _vp, _vp, _vp]),
    "gxa_block_fwd": (_i, [_dp, _bp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gxa_block_bwd": (_i, [_dp, _bp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _bp, _vp, _vp]),
    "gxa_block_kv_cache_bytes": (_sz, [_dp]),
    "gxa_block_kv_prefill": (_i, [_dp, _bp, _vp, _vp, _vp, _vp]),
    "gxa_block_decode_workspace_bytes": (_sz, [_dp]),
    "gxa_block_decode": (_i, [_dp, _bp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gxa_layernorm_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _vp]),
    "gxa_layernorm_pe_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _vp]),
    "gxa_sum3_f32": (_i, [_vp, _vp, _vp, _vp, _sz, _vp]),
    "gxa_layernorm_bwd": (_i, [_vp] * 9 + [_i, _i, _vp]),
    "gxa_linear_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "gxa_linear_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp]),
    "gxa_linear_bwd": (_i, [_vp] * 6 + [_i, _i, _i, _i, _vp, _vp]),
    "gxa_linear_few_rows_bf16": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _i, _vp, _i, _vp, _vp, _i, _vp, _vp]),
    "gxa_linear_fwd_static": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "gxa_gemm_bf16": (_i, [_vp, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp, _i, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _i,
                           _vp]),
    "gxa_attn_fwd": (_i, [_vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _vp]),
    "gxa_attn_fwd_dropout": (_i, [_vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, C.c_float, C.c_uint64,
                                  _vp]),
    "gxa_attn_bwd_workspace_bytes": (_sz, [_i, _i, _i]),
    "gxa_attn_bwd": (_i, [_vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _i,
                          _i, _vp]),
    "gxa_attn_bwd_dropout": (_i, [_vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp, _i, _vp, _i,
                                  _i, _i, _i, C.c_float, C.c_uint64, _vp]),
}

_lib = None


def lib() -> C.CDLL:
    """Load (once) and return the library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m mocov2_whisper_flamingo_b200.build` "
                "(there is no CPU/PyTorch fallback for the gated cross-attention path)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(code: int, what: str) -> None:
    if code != 0:
        detail = lib().gxa_last_cuda_error().decode() if code in (-3, -4) else ""
        raise RuntimeError(f"{what} failed: {_ERR.get(code, code)} {detail}".rstrip())


def launch_count() -> int:
    return int(lib().gxa_launch_count())


def set_sm_limit(n_sms: int) -> int:
    """SMs the persistent grids may assume (0 = all); returns the previous limit (include/gxa.h)."""
    return int(lib().gxa_set_sm_limit(int(n_sms)))
