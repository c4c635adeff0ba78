"""ctypes binding of the C ABI declared in include/adaface_b200.h.

The product path fails loudly when the CUDA library is missing or unusable: there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

_LIB = None


class AdafaceB200Error(RuntimeError):
    pass


class GemmArgs(C.Structure):
    _fields_ = [("a", C.c_void_p), ("w", C.c_void_p), ("out", C.c_void_p), ("bias", C.c_void_p),
                ("residual", C.c_void_p), ("m", C.c_int64), ("n", C.c_int64), ("k", C.c_int64),
                ("lda", C.c_int64), ("ldw", C.c_int64), ("ldo", C.c_int64), ("ldr", C.c_int64),
                ("epilogue", C.c_int), ("out_dtype", C.c_int), ("residual_dtype", C.c_int), ("hw", C.c_int),
                ("row_stats_out", C.c_void_p), ("ln_stats", C.c_void_p), ("ln_colsum", C.c_void_p),
                ("ln_parts", C.c_int), ("ln_eps", C.c_float), ("conv_h", C.c_int), ("conv_w", C.c_int), ("conv_fold", C.c_int), ("a2", C.c_void_p), ("k_split", C.c_int64), ("lda2", C.c_int64), ("conv_stride", C.c_int)]


class AdaPrepItem(C.Structure):
    _fields_ = [("to_k_weight", C.c_void_p), ("fg_q_emb", C.c_void_p), ("bg_q_emb", C.c_void_p),
                ("lora_to_fg_q", C.c_void_p), ("lora_to_bg_q", C.c_void_p), ("lora_to_k", C.c_void_p),
                ("pool_q", C.c_void_p), ("ws", C.c_void_p), ("channels", C.c_int), ("reserved", C.c_int)]


class AttnArgs(C.Structure):
    _fields_ = [("q", C.c_void_p), ("k", C.c_void_p), ("v", C.c_void_p), ("out", C.c_void_p),
                ("key_mask", C.c_void_p), ("batch", C.c_int), ("heads", C.c_int), ("head_dim", C.c_int),
                ("q_len", C.c_int), ("kv_len", C.c_int), ("ldq", C.c_int64), ("ldk", C.c_int64),
                ("ldv", C.c_int64), ("ldo", C.c_int64), ("scale", C.c_float), ("lse", C.c_void_p)]


EPI_STORE, EPI_GEGLU, EPI_NCHW = 0, 1, 2
DT_BF16, DT_F32 = 0, 1

_P, _I, _I64, _F = C.c_void_p, C.c_int, C.c_int64, C.c_float
_SIGNATURES = {
    "afb_last_error": (C.c_char_p, []),
    "afb_abi_version": (_I, []),
    "afb_header_digest": (C.c_char_p, []),
    "afb_launch_count": (_I64, []),
    "afb_cfg_ddim_step": (_I, [_P, _P, _P, _P, _P, _P, _I64, _F, _F, _F, _F, _F, _F, _P]),
    "afb_groupnorm_ws_floats": (_I64, [_I, _I, _I]),
    "afb_groupnorm_tokens": (_I, [_P, _I, _P, _P, _P, _P, _I, _I, _I, _I, _F, _I, _P, _I, _P]),
    "afb_groupnorm_tokens_cat": (_I, [_P, _I, _P, _I, _P, _P, _P, _P, _I, _I, _I, _I, _F, _I, _P, _I, _P]),
    "afb_groupnorm_static_ws_max_floats": (_I64, []),
    "afb_groupnorm_tokens_static_ws": (_I, [_P, _I, _P, _P, _P, _I, _I, _I, _I, _F, _I, _P, _I, _P]),
    "afb_add_bias_residual": (_I, [_P, _P, _P, _P, _I64, _I, _P]),
    "afb_layernorm_bf16": (_I, [_P, _P, _P, _P, _I64, _I, _F, _P]),
    "afb_gemm_bf16": (_I, [C.POINTER(GemmArgs), _P]),
    "afb_gemm_stats_parts": (_I, [_I64]),
    "afb_im2col3x3_c4": (_I, [_P, _P, _I, _I, _I, _P]),
    "afb_conv3x3_few_out": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "afb_attention_fwd": (_I, [C.POINTER(AttnArgs), _P]),
    "afb_cross_scores": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I64, _I64, _F, _P]),
    "afb_conv_attn_rows": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I64, _I64, _I, _I, _F, _P]),
    "afb_softmax_pv": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _I64, _I64, _P]),
    "afb_ada_pool_ws_floats": (_I64, [_I, _I, _I]),
    "afb_ada_pool_prep_ws_floats": (_I64, [_I, _I]),
    "afb_ada_pool_stats_offset": (_I64, [_I, _I, _I]),
    "afb_ada_pool_prep": (_I, [_P, _I, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P]),
    "afb_ada_pool_prep_batched": (_I, [_P, _I, _I, _I, _I, _P]),
    "afb_ada_pool": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _F, C.c_uint32, _P]),
    "afb_ada_head": (_I, [_P, _P, _P, _I, _P, _P, _P, _P, _P, _P, _P, _I64, _I64, _I, _I, _I, _I, _I, _I, _P]),
    "afb_ada_head_pooled": (_I, [_P, _I, _P, _P, _P, _P, _I, _P, _P, _P, _P, _P, _P, _P, _I64, _I64, _I, _I, _I, _I, _I, _I, _P]),
    "afb_static_layerwise_embedding": (_I, [_P, _P, _P, _I, _P, _P, _P, _I, _I, _I, _I, _P]),
    "afb_first_token_pos": (_I, [_P, _I64, _P, _I, _I, _P]),
    "afb_splice_static_embedding": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "afb_geglu_bwd": (_I, [_P, _P, _P, _I64, _I, _P]),
    "afb_layernorm_bwd": (_I, [_P, _P, _P, _P, _I64, _I, _F, _P]),
    "afb_attention_bwd_short": (_I, [C.POINTER(AttnArgs), _P, _P, _P, _P, _I64, _I64, _P]),
    "afb_ada_head_bwd": (_I, [_P, _P, _P, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _P]),
    "afb_ada_pool_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _F, C.c_uint32, _P]),
    "afb_ada_pool_prep_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _P]),
    "afb_attention_bwd": (_I, [C.POINTER(AttnArgs), _P, _P, _P, _P, _P, _I64, _P]),
    "afb_groupnorm_tokens_bwd": (_I, [_P, _P, _P, _P, _P, _I64, _I, _P, _P, _I, _I, _I, _I, _F, _P]),
    "afb_ada_pool_bwd_features": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _F, C.c_uint32, _P]),
    "afb_cast_f32_to_bf16": (_I, [_P, _P, _I64, _P]),
    "afb_cast_bf16_to_f32": (_I, [_P, _P, _I64, _P]),
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def lib_path() -> str:
    return _build.LIB_PATH


def load():
    """Loads (building first if sources changed and nvcc is present) the shared library."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not _build.is_current():
        try:
            _build.build()
        except Exception as exc:  # no nvcc on this machine: accept a prebuilt library, else fail loudly
            if not os.path.exists(path):
                raise AdafaceB200Error(
                    f"adaface_b200: CUDA library {path} is missing and could not be built ({exc}). "
                    "There is no CPU fallback for this path.") from exc
    try:
        lib = C.CDLL(path)
    except OSError as exc:
        raise AdafaceB200Error(f"adaface_b200: cannot load {path}: {exc}. There is no CPU fallback.") from exc
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = ABI mismatch, fail loudly
        fn.restype = res
        fn.argtypes = args
    # a prebuilt library that could not be rebuilt (no nvcc here) must at least have been built against THIS header:
    # a changed afb_gemm_args / afb_attn_args layout or argument order would otherwise corrupt kernel arguments silently
    built_against = (lib.afb_header_digest() or b"").decode()
    if built_against != _build.header_digest():
        raise AdafaceB200Error(
            f"adaface_b200: {path} was built against another revision of include/adaface_b200.h "
            f"({built_against[:12]}... vs {_build.header_digest()[:12]}...) and could not be rebuilt here; "
            "run `python -m adaface_b200.build` where nvcc is available")
    _LIB = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().afb_last_error()
        raise AdafaceB200Error(f"{what or 'adaface_b200'} failed (code {rc}): {msg.decode() if msg else ''}")


def launch_count() -> int:
    return int(load().afb_launch_count())
