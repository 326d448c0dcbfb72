"""Random-init decoder weights in the flat layout the C-ABI takes (include/b200score.h).

All BASELINE configs are random-init (no checkpoints offline): linears and
embeddings ~ N(0, 0.02) as HF ``_init_weights`` does (HF:models/falcon/modeling_falcon.py:649-656).
For parity tests LayerNorm affine parameters and biases are drawn randomly too, so those code
paths carry signal.  The same tensors are loaded into the HF model that serves as the oracle
(``oracle/hf_models.py``, test infrastructure), so both sides see identical (bf16-representable when asked) numbers.

Slot order (``B2S_W_*`` in the header):
  global:    0 wte[V,d]  1 lnf_g[d]  2 lnf_b[d]  3 lm_head[V,d]  4 rope_cos[max_pos,rd/2]  5 rope_sin
             6 wpe[max_pos+pos_offset,d] (learned positions)  7 emb_ln_g[d]  8 emb_ln_b[d]  9 alibi_slopes[n_head]
  per layer: base = 10 + 16*l
             +0 ln1_g +1 ln1_b +2 ln2_g +3 ln2_b +4 w_qkv[qkv,d] +5 b_qkv +6 w_dense[d,nh*hd] +7 b_dense
             +8 w_up[ff,d] ([2ff,d] = gate rows then up rows for SwiGLU) +9 b_up +10 w_down[d,ff] +11 b_down   (+12..15 reserved)
RMSNorm models have no *_b vectors (null slots).
Matrices are stored in the model's weight dtype (bf16 or fp32); vectors are always fp32.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from .shapes import DecoderShape, NORM_RMS, POS_LEARNED, POS_ALIBI

N_GLOBAL = 10
N_PER_LAYER = 16


def rope_tables(shape: DecoderShape, device="cpu"):
    """cos/sin of HF's rotary embedding (HF falcon:154-165): fp32 inv_freq x fp32 position, first
    rotary_dim/2 columns (the table HF builds is cat(freqs, freqs))."""
    rd = shape.rotary_dim
    inv_freq = 1.0 / (shape.rope_theta ** (torch.arange(0, rd, 2, dtype=torch.int64).float() / rd))
    pos = torch.arange(shape.max_pos, dtype=torch.float32)
    freqs = pos[:, None] * inv_freq[None, :]
    return freqs.cos().contiguous().to(device), freqs.sin().contiguous().to(device)


def slot_names(shape: DecoderShape) -> List[Optional[str]]:
    names: List[Optional[str]] = ["wte", "lnf_g", "lnf_b", "lm_head", "rope_cos", "rope_sin", "wpe", "emb_ln_g", "emb_ln_b",
                                  "alibi_slopes"]
    per = ["ln1_g", "ln1_b", "ln2_g", "ln2_b", "w_qkv", "b_qkv", "w_dense", "b_dense",
           "w_up", "b_up", "w_down", "b_down", None, None, None, None]
    for l in range(shape.n_layer):
        for p in per:
            names.append(None if p is None else f"l{l}.{p}")
    return names


def is_matrix(name: str) -> bool:
    leaf = name.split(".")[-1]
    return leaf in ("wte", "wpe", "lm_head", "w_qkv", "w_dense", "w_up", "w_down")


def alibi_slopes(n_head: int) -> torch.Tensor:
    """Per-head ALiBi slopes exactly as HF:models/bloom/modeling_bloom.py build_alibi_tensor computes them (fp32):
    powers of 2^(-8/n) for the closest power of two n <= n_head, then odd powers of the half-step base for the rest."""
    import math
    cp2 = 2 ** math.floor(math.log2(n_head))
    base = torch.tensor(2 ** (-(2 ** -(math.log2(cp2) - 3))), dtype=torch.float32)
    slopes = torch.pow(base, torch.arange(1, 1 + cp2, dtype=torch.int32))
    if cp2 != n_head:
        extra_base = torch.tensor(2 ** (-(2 ** -(math.log2(2 * cp2) - 3))), dtype=torch.float32)
        n_rem = min(cp2, n_head - cp2)
        slopes = torch.cat([slopes, torch.pow(extra_base, torch.arange(1, 1 + 2 * n_rem, 2, dtype=torch.int32))], dim=0)
    return slopes.to(torch.float32).contiguous()


def init_weights(shape: DecoderShape, seed: int = 0, device: str = "cpu",
                 randomize_affine: bool = True, bf16_representable: bool = True,
                 std: float = 0.02) -> Dict[str, torch.Tensor]:
    """fp32 master copy.  With ``bf16_representable`` every matrix entry is rounded to bf16
    (then widened back) so a bf16 device copy and the fp32 oracle hold identical values."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    d, ff, V = shape.d_model, shape.d_ff, shape.vocab

    def mat(r, c):
        w = torch.empty(r, c, dtype=torch.float32, device=device).normal_(0.0, std, generator=g)
        if bf16_representable:
            w = w.to(torch.bfloat16).to(torch.float32)
        return w

    def vec(n, base, spread):
        if not randomize_affine:
            return torch.full((n,), float(base), dtype=torch.float32, device=device)
        return base + spread * torch.empty(n, dtype=torch.float32, device=device).normal_(0.0, 1.0, generator=g)

    W: Dict[str, torch.Tensor] = {}
    W["wte"] = mat(V, d)
    rms = shape.norm_type == NORM_RMS
    W["lnf_g"] = vec(d, 1.0, 0.1)
    if not rms:
        W["lnf_b"] = vec(d, 0.0, 0.05)
    W["lm_head"] = W["wte"] if shape.tied_head else mat(V, d)
    if shape.pos_type == POS_LEARNED:
        W["wpe"] = mat(shape.max_pos + shape.pos_offset, d)
    if shape.emb_norm:
        W["emb_ln_g"] = vec(d, 1.0, 0.1)
        W["emb_ln_b"] = vec(d, 0.0, 0.05)
    if shape.pos_type == POS_ALIBI:
        W["alibi_slopes"] = alibi_slopes(shape.n_head).to(device)
    for l in range(shape.n_layer):
        p = f"l{l}."
        W[p + "ln1_g"] = vec(d, 1.0, 0.1)
        if not rms:
            W[p + "ln1_b"] = vec(d, 0.0, 0.05)
        if shape.n_ln_per_layer == 2:
            W[p + "ln2_g"] = vec(d, 1.0, 0.1)
            if not rms:
                W[p + "ln2_b"] = vec(d, 0.0, 0.05)
        W[p + "w_qkv"] = mat(shape.qkv_dim, d)
        W[p + "w_dense"] = mat(d, shape.n_head * shape.head_dim)
        W[p + "w_up"] = mat(shape.up_dim, d)
        W[p + "w_down"] = mat(d, ff)
        if shape.qkv_bias and not shape.has_bias:
            W[p + "b_qkv"] = 0.0 + 0.2 * torch.empty(shape.qkv_dim, dtype=torch.float32, device=device).normal_(0.0, 1.0, generator=g)
        if shape.has_bias:
            W[p + "b_qkv"] = vec(shape.qkv_dim, 0.0, 0.02)
            W[p + "b_dense"] = vec(d, 0.0, 0.02)
            W[p + "b_up"] = vec(ff, 0.0, 0.02)
            W[p + "b_down"] = vec(d, 0.0, 0.02)
    return W


def init_weights_device(shape: DecoderShape, seed: int, device, weight_dtype=torch.bfloat16,
                        randomize_affine: bool = False, std: float = 0.02) -> Dict[str, torch.Tensor]:
    """Full-size random init directly on the GPU (replicated on every rank from the same seed).
    Matrices are produced layer by layer in the target dtype so that a 7B model never holds an
    fp32 copy."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    d, ff, V = shape.d_model, shape.d_ff, shape.vocab

    def mat(r, c):
        w = torch.empty(r, c, dtype=torch.float32, device=device).normal_(0.0, std, generator=g)
        return w.to(weight_dtype)

    def vec(n, base, spread):
        if not randomize_affine:
            return torch.full((n,), float(base), dtype=torch.float32, device=device)
        return base + spread * torch.empty(n, dtype=torch.float32, device=device).normal_(0.0, 1.0, generator=g)

    W: Dict[str, torch.Tensor] = {}
    W["wte"] = mat(V, d)
    rms = shape.norm_type == NORM_RMS
    W["lnf_g"] = vec(d, 1.0, 0.1)
    if not rms:
        W["lnf_b"] = vec(d, 0.0, 0.05)
    W["lm_head"] = W["wte"] if shape.tied_head else mat(V, d)
    if shape.pos_type == POS_LEARNED:
        W["wpe"] = mat(shape.max_pos + shape.pos_offset, d)
    if shape.emb_norm:
        W["emb_ln_g"] = vec(d, 1.0, 0.1)
        W["emb_ln_b"] = vec(d, 0.0, 0.05)
    if shape.pos_type == POS_ALIBI:
        W["alibi_slopes"] = alibi_slopes(shape.n_head).to(device)
    for l in range(shape.n_layer):
        p = f"l{l}."
        W[p + "ln1_g"] = vec(d, 1.0, 0.1)
        if not rms:
            W[p + "ln1_b"] = vec(d, 0.0, 0.05)
        if shape.n_ln_per_layer == 2:
            W[p + "ln2_g"] = vec(d, 1.0, 0.1)
            if not rms:
                W[p + "ln2_b"] = vec(d, 0.0, 0.05)
        W[p + "w_qkv"] = mat(shape.qkv_dim, d)
        W[p + "w_dense"] = mat(d, shape.n_head * shape.head_dim)
        W[p + "w_up"] = mat(shape.up_dim, d)
        W[p + "w_down"] = mat(d, ff)
        if shape.qkv_bias and not shape.has_bias:
            W[p + "b_qkv"] = 0.0 + 0.2 * torch.empty(shape.qkv_dim, dtype=torch.float32, device=device).normal_(0.0, 1.0, generator=g)
        if shape.has_bias:
            W[p + "b_qkv"] = vec(shape.qkv_dim, 0.0, 0.02)
            W[p + "b_dense"] = vec(d, 0.0, 0.02)
            W[p + "b_up"] = vec(ff, 0.0, 0.02)
            W[p + "b_down"] = vec(d, 0.0, 0.02)
    return W


def to_device(W: Dict[str, torch.Tensor], device, weight_dtype) -> Dict[str, torch.Tensor]:
    out = {}
    for k, v in W.items():
        if k == "lm_head" and v is W.get("wte"):
            continue
        out[k] = v.to(device=device, dtype=weight_dtype if is_matrix(k) else torch.float32).contiguous()
    if "lm_head" not in out:
        out["lm_head"] = out["wte"]
    return out


def pointer_table(shape: DecoderShape, Wdev: Dict[str, torch.Tensor]):
    """ctypes array of device pointers in slot order (null where the model has no such tensor)."""
    import ctypes as C
    names = slot_names(shape)
    arr = (C.c_void_p * len(names))()
    for i, n in enumerate(names):
        t = Wdev.get(n) if n is not None else None
        arr[i] = None if t is None else t.data_ptr()
    return arr
