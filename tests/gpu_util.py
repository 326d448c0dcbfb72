"""Helpers shared by the -m gpu parity tests (all calls go through the C-ABI library)."""
import ctypes as C
import json
import os

import numpy as np
import torch

from conftest import GOLDEN
from llm_interpretation_replication_b200 import _lib, shapes, weights
from llm_interpretation_replication_b200.scorer import Scorer
from llm_interpretation_replication_b200.tokenizer import HashTokenizer

DEV = "cuda:0"


def stream():
    return _lib.current_stream()


def gemm_bf16(A, W, bias, epi, out=None):
    lib = _lib.load()
    M, K = A.shape
    N = W.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32 if epi in (2, 4) else torch.bfloat16, device=A.device)
    _lib.check(lib.b2s_gemm_bf16(_lib.ptr(A), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(out), M, N, K, epi, stream()), "gemm_bf16")
    return out


def gemm_f16(A, W, bias, epi, out=None):
    """fp16 A x fp16 W (the FP16 mode's GEMM)."""
    lib = _lib.load()
    M, K = A.shape
    N = W.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32 if epi in (2, 4) else torch.float16, device=A.device)
    _lib.check(lib.b2s_gemm_f16(_lib.ptr(A), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(out), M, N, K, epi, stream()), "gemm_f16")
    return out


def gemm_f32(A, W, bias, epi, out=None):
    lib = _lib.load()
    M, K = A.shape
    N = W.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.float32, device=A.device)
    _lib.check(lib.b2s_gemm_f32(_lib.ptr(A), _lib.ptr(W), _lib.ptr(bias), _lib.ptr(out), M, N, K, epi, stream()), "gemm_f32")
    return out


def lm_head(hidden, W, cand):
    lib = _lib.load()
    M, d = hidden.shape
    V = W.shape[0]
    nbytes = lib.b2s_lm_head_scratch_bytes(M, V)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=hidden.device)
    stats = torch.empty((M, 16), dtype=torch.float32, device=hidden.device)
    _lib.check(lib.b2s_lm_head_bf16(_lib.ptr(hidden), _lib.ptr(W), _lib.ptr(cand), M, V, d, _lib.ptr(stats),
                                    _lib.ptr(scratch), nbytes, stream()), "lm_head")
    return stats


def make_scorer(shp, W, mode, **kw):
    dt = torch.float32 if mode == shapes.MODE_FP32_CHECK else torch.bfloat16
    return Scorer(shp, weights.to_device(W, DEV, dt), mode=mode, device=DEV, **kw)


def golden_cases():
    return json.load(open(os.path.join(GOLDEN, "scorer_golden.json")))["cases"]


def tok_for(shp, row=None):
    tok = HashTokenizer(shp.vocab, eos_token_id=shp.eos_id).apply_pad_default()
    if row is not None and row.get("cand_override") is not None:
        tok.overrides = {" Yes": row["cand_override"][0], " No": row["cand_override"][1]}
    return tok
