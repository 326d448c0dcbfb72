"""ctypes binding of libb200score.so (include/b200score.h).  There is no fallback: if the shared
library is missing or a call fails, an exception is raised with the library's own message."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200score.so")

ABI_VERSION = 4


class B2SError(RuntimeError):
    pass


class B2SRangeError(B2SError):
    """B2S_ERANGE: the call completed, but some prompts produced non-finite logits (their rows are NaN with found = -1)."""


E_RANGE = -5


class ModelDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("arch", C.c_int32),
        ("n_layer", C.c_int32), ("d_model", C.c_int32), ("n_head", C.c_int32), ("n_kv_head", C.c_int32),
        ("head_dim", C.c_int32), ("d_ff", C.c_int32), ("vocab", C.c_int32), ("rotary_dim", C.c_int32),
        ("norm_eps", C.c_float),
        ("parallel_residual", C.c_int32), ("n_ln_per_layer", C.c_int32), ("has_bias", C.c_int32),
        ("qkv_layout", C.c_int32), ("max_pos", C.c_int32), ("compute_mode", C.c_int32),
        ("norm_type", C.c_int32), ("mlp_type", C.c_int32),
        ("pos_type", C.c_int32), ("pos_offset", C.c_int32), ("act_type", C.c_int32), ("emb_norm", C.c_int32),
    ]


class EncDecDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("n_enc_layer", C.c_int32), ("n_dec_layer", C.c_int32), ("d_model", C.c_int32),
        ("n_head", C.c_int32), ("head_dim", C.c_int32), ("d_ff", C.c_int32), ("vocab", C.c_int32),
        ("norm_eps", C.c_float), ("ff_type", C.c_int32), ("max_pos", C.c_int32), ("compute_mode", C.c_int32),
        ("decoder_start_id", C.c_int32),
    ]


# every symbol include/b200score.h declares: name -> (restype, argtypes)
_vp, _i32, _i64, _u32, _u64, _sz, _dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_uint32, C.c_uint64, C.c_size_t, C.c_double
SIGNATURES = {
    "b2s_last_error": (C.c_char_p, []),
    "b2s_abi_version": (C.c_int, []),
    "b2s_model_create": (C.c_int, [C.POINTER(ModelDesc), C.POINTER(_vp), _i32, _i32, C.POINTER(_vp)]),
    "b2s_model_destroy": (None, [_vp]),
    "b2s_workspace_bytes": (_sz, [_vp, _i64, _i32, _i32]),
    "b2s_score": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
    "b2s_score_host": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
    "b2s_score_complete": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
    "b2s_score_complete_host": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
    "b2s_encdec_create": (C.c_int, [C.POINTER(EncDecDesc), C.POINTER(_vp), _i32, _i32, C.POINTER(_vp)]),
    "b2s_encdec_destroy": (None, [_vp]),
    "b2s_encdec_workspace_bytes": (_sz, [_vp, _i64, _i32, _i32]),
    "b2s_encdec_score": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
    "b2s_encdec_score_host": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
    "b2s_int8_fakequant_f16": (C.c_int, [_vp, _i32, _i32, _vp, _vp, _i32, _vp, C.c_float, _vp]),
    "b2s_launch_count": (_i64, [_i32]),
    "b2s_set_option": (C.c_int, [C.c_char_p, _i32]),
    "b2s_profile_gemm": (C.c_int, [_i32]),
    "b2s_profile_gemm_read": (C.c_int, [C.POINTER(C.c_double), C.POINTER(_i64), C.POINTER(C.c_double)]),
    "b2s_gemm_bf16": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "b2s_gemm_f16": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "b2s_gemm_f32": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp]),
    "b2s_attention_bf16": (C.c_int, [_vp] * 8 + [_i32] * 12 + [_vp]),
    "b2s_lm_head_scratch_bytes": (_sz, [_i32, _i32]),
    "b2s_lm_head_bf16": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _sz, _vp]),
    "b2s_lm_head_static_capacity": (_sz, []),
    "b2s_lm_head_bf16_static": (C.c_int, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    "b2s_mt19937_seed": (C.c_int, [_u32, _vp, _vp]),
    "b2s_mt19937_choice": (C.c_int, [_vp, _i32, _i64, _vp, _vp]),
    "b2s_bootstrap_resample": (C.c_int, [_vp, _i32, _i32, _i32, _i64, _vp, _u64, _vp, _vp]),
    "b2s_ci_simulate": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i32, _i32, _u32, _i32, _vp, _vp, _vp, _vp]),
    "b2s_corr_bootstrap": (C.c_int, [_vp, _i32, _i32, _vp, _i32, _vp, _i32, _i32, _vp, _vp, _vp]),
    "b2s_corr_bootstrap_ex": (C.c_int, [_vp, _i32, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _vp, _vp, _vp]),
    "b2s_philox_indices": (C.c_int, [_u64, _i32, _i64, _vp, _vp]),
    "b2s_agreement_bootstrap": (C.c_int, [_vp, _vp, _i32, _vp, _i64, _i32, _vp, _vp]),
    "b2s_bootstrap_scratch_bytes": (_sz, [_i64]),
    "b2s_bootstrap_summary": (C.c_int, [_vp, _i64, _i32, _dbl, _dbl, _dbl, _vp, _vp, _vp, _sz, _vp]),
}

_lib = None


def load():
    """Load the library (once).  Raises B2SError if it has not been built: there is no CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B2SError(f"{LIB_PATH} not found: build it with `python -m llm_interpretation_replication_b200.build` "
                       "(there is no CPU or PyTorch fallback for this path)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    got = lib.b2s_abi_version()
    if got != ABI_VERSION:
        raise B2SError(f"libb200score ABI {got}, Python binding expects {ABI_VERSION}")
    _lib = lib
    return lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().b2s_last_error().decode("utf-8", "replace")
        raise (B2SRangeError if rc == E_RANGE else B2SError)(f"{what}: error {rc}: {msg}")


def ptr(t):
    """data_ptr of a torch tensor (or None) as c_void_p."""
    return None if t is None else C.c_void_p(t.data_ptr())


def current_stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
