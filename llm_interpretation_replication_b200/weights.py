"""Random-init decoder weights in the flat layout the C-ABI takes (include/b200score.h).

All BASELINE configs are random-init (no checkpoints offline): linears and
embeddings ~ N(0, 0.02) as HF ``_init_weights`` does (HF:models/falcon/modeling_falcon.py:649-656).
For parity tests LayerNorm affine parameters and biases are drawn randomly too, so those code
paths carry signal.  The same tensors are loaded into the HF model that serves as the oracle
(``to_hf_state_dict``), so both sides see identical (bf16-representable when asked) numbers.

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

from .shapes import (DecoderShape, ARCH_FALCON, ARCH_GPT_NEOX, ARCH_LLAMA, ARCH_OPT, ARCH_BLOOM, NORM_RMS, POS_LEARNED,
                     POS_ALIBI)

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


# ---- HF oracle side ---------------------------------------------------------------------
def hf_config(shape: DecoderShape):
    """HF config equivalent to ``shape`` (used by the oracle and the golden generator only)."""
    if shape.arch == ARCH_FALCON:
        from transformers import FalconConfig
        return FalconConfig(vocab_size=shape.vocab, hidden_size=shape.d_model,
                            num_hidden_layers=shape.n_layer, num_attention_heads=shape.n_head,
                            multi_query=True, parallel_attn=True, bias=False, alibi=False,
                            new_decoder_architecture=False, layer_norm_epsilon=shape.norm_eps,
                            bos_token_id=shape.eos_id, eos_token_id=shape.eos_id,
                            max_position_embeddings=shape.max_pos, tie_word_embeddings=True,
                            hidden_dropout=0.0, attention_dropout=0.0)
    if shape.arch == ARCH_GPT_NEOX:
        from transformers import GPTNeoXConfig
        return GPTNeoXConfig(vocab_size=shape.vocab, hidden_size=shape.d_model,
                             num_hidden_layers=shape.n_layer, num_attention_heads=shape.n_head,
                             intermediate_size=shape.d_ff, hidden_act="gelu",
                             rotary_pct=shape.rotary_dim / shape.head_dim, rotary_emb_base=shape.rope_theta,
                             layer_norm_eps=shape.norm_eps, use_parallel_residual=shape.parallel_residual,
                             bos_token_id=shape.eos_id, eos_token_id=shape.eos_id,
                             max_position_embeddings=shape.max_pos, tie_word_embeddings=False,
                             attention_bias=True, hidden_dropout=0.0, attention_dropout=0.0)
    if shape.arch == ARCH_LLAMA:
        from transformers import LlamaConfig
        return LlamaConfig(vocab_size=shape.vocab, hidden_size=shape.d_model, intermediate_size=shape.d_ff,
                           num_hidden_layers=shape.n_layer, num_attention_heads=shape.n_head,
                           num_key_value_heads=shape.n_kv_head, head_dim=shape.head_dim, hidden_act="silu",
                           max_position_embeddings=shape.max_pos, rms_norm_eps=shape.norm_eps, rope_theta=shape.rope_theta,
                           attention_bias=False, mlp_bias=False, tie_word_embeddings=shape.tied_head,
                           bos_token_id=1, eos_token_id=shape.eos_id, attention_dropout=0.0)
    if shape.arch == ARCH_OPT:
        from transformers import OPTConfig
        return OPTConfig(vocab_size=shape.vocab, hidden_size=shape.d_model, num_hidden_layers=shape.n_layer,
                         ffn_dim=shape.d_ff, num_attention_heads=shape.n_head, max_position_embeddings=shape.max_pos,
                         do_layer_norm_before=True, word_embed_proj_dim=shape.d_model, activation_function="relu",
                         dropout=0.0, attention_dropout=0.0, layerdrop=0.0, enable_bias=True,
                         layer_norm_elementwise_affine=True, bos_token_id=shape.eos_id, eos_token_id=shape.eos_id,
                         pad_token_id=1, tie_word_embeddings=True)
    if shape.arch == ARCH_BLOOM:
        from transformers import BloomConfig
        return BloomConfig(vocab_size=shape.vocab, hidden_size=shape.d_model, n_layer=shape.n_layer, n_head=shape.n_head,
                           layer_norm_epsilon=shape.norm_eps, apply_residual_connection_post_layernorm=False,
                           hidden_dropout=0.0, attention_dropout=0.0, bos_token_id=1, eos_token_id=shape.eos_id,
                           pad_token_id=3, tie_word_embeddings=True)
    raise ValueError(shape.arch)


def to_hf_state_dict(shape: DecoderShape, W: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    sd = {}
    if shape.arch == ARCH_FALCON:
        sd["transformer.word_embeddings.weight"] = W["wte"]
        sd["transformer.ln_f.weight"] = W["lnf_g"]
        sd["transformer.ln_f.bias"] = W["lnf_b"]
        sd["lm_head.weight"] = W["lm_head"]
        for l in range(shape.n_layer):
            p, h = f"l{l}.", f"transformer.h.{l}."
            sd[h + "input_layernorm.weight"] = W[p + "ln1_g"]
            sd[h + "input_layernorm.bias"] = W[p + "ln1_b"]
            sd[h + "self_attention.query_key_value.weight"] = W[p + "w_qkv"]
            sd[h + "self_attention.dense.weight"] = W[p + "w_dense"]
            sd[h + "mlp.dense_h_to_4h.weight"] = W[p + "w_up"]
            sd[h + "mlp.dense_4h_to_h.weight"] = W[p + "w_down"]
    elif shape.arch == ARCH_LLAMA:
        # HF:models/llama/modeling_llama.py: separate q/k/v and gate/up projections; the flat layout fuses them by rows
        nq, nk, ff = shape.n_head * shape.head_dim, shape.n_kv_head * shape.head_dim, shape.d_ff
        sd["model.embed_tokens.weight"] = W["wte"]
        sd["model.norm.weight"] = W["lnf_g"]
        sd["lm_head.weight"] = W["lm_head"]
        for l in range(shape.n_layer):
            p, h = f"l{l}.", f"model.layers.{l}."
            sd[h + "input_layernorm.weight"] = W[p + "ln1_g"]
            sd[h + "post_attention_layernorm.weight"] = W[p + "ln2_g"]
            sd[h + "self_attn.q_proj.weight"] = W[p + "w_qkv"][:nq]
            sd[h + "self_attn.k_proj.weight"] = W[p + "w_qkv"][nq:nq + nk]
            sd[h + "self_attn.v_proj.weight"] = W[p + "w_qkv"][nq + nk:]
            sd[h + "self_attn.o_proj.weight"] = W[p + "w_dense"]
            sd[h + "mlp.gate_proj.weight"] = W[p + "w_up"][:ff]
            sd[h + "mlp.up_proj.weight"] = W[p + "w_up"][ff:]
            sd[h + "mlp.down_proj.weight"] = W[p + "w_down"]
    elif shape.arch == ARCH_OPT:
        # HF:models/opt/modeling_opt.py: separate q/k/v projections with biases; fused by rows in the flat layout
        nq = shape.n_head * shape.head_dim
        b = "model.decoder."
        sd[b + "embed_tokens.weight"] = W["wte"]
        sd[b + "embed_positions.weight"] = W["wpe"]
        sd[b + "final_layer_norm.weight"] = W["lnf_g"]
        sd[b + "final_layer_norm.bias"] = W["lnf_b"]
        sd["lm_head.weight"] = W["lm_head"]
        for l in range(shape.n_layer):
            p, h = f"l{l}.", f"{b}layers.{l}."
            sd[h + "self_attn_layer_norm.weight"] = W[p + "ln1_g"]
            sd[h + "self_attn_layer_norm.bias"] = W[p + "ln1_b"]
            sd[h + "final_layer_norm.weight"] = W[p + "ln2_g"]
            sd[h + "final_layer_norm.bias"] = W[p + "ln2_b"]
            for i, n in enumerate(("q_proj", "k_proj", "v_proj")):
                sd[h + f"self_attn.{n}.weight"] = W[p + "w_qkv"][i * nq:(i + 1) * nq]
                sd[h + f"self_attn.{n}.bias"] = W[p + "b_qkv"][i * nq:(i + 1) * nq]
            sd[h + "self_attn.out_proj.weight"] = W[p + "w_dense"]
            sd[h + "self_attn.out_proj.bias"] = W[p + "b_dense"]
            sd[h + "fc1.weight"] = W[p + "w_up"]
            sd[h + "fc1.bias"] = W[p + "b_up"]
            sd[h + "fc2.weight"] = W[p + "w_down"]
            sd[h + "fc2.bias"] = W[p + "b_down"]
    elif shape.arch == ARCH_BLOOM:
        # HF:models/bloom/modeling_bloom.py: fused query_key_value with per-head [q, k, v] rows
        sd["transformer.word_embeddings.weight"] = W["wte"]
        sd["transformer.word_embeddings_layernorm.weight"] = W["emb_ln_g"]
        sd["transformer.word_embeddings_layernorm.bias"] = W["emb_ln_b"]
        sd["transformer.ln_f.weight"] = W["lnf_g"]
        sd["transformer.ln_f.bias"] = W["lnf_b"]
        sd["lm_head.weight"] = W["lm_head"]
        for l in range(shape.n_layer):
            p, h = f"l{l}.", f"transformer.h.{l}."
            sd[h + "input_layernorm.weight"] = W[p + "ln1_g"]
            sd[h + "input_layernorm.bias"] = W[p + "ln1_b"]
            sd[h + "post_attention_layernorm.weight"] = W[p + "ln2_g"]
            sd[h + "post_attention_layernorm.bias"] = W[p + "ln2_b"]
            sd[h + "self_attention.query_key_value.weight"] = W[p + "w_qkv"]
            sd[h + "self_attention.query_key_value.bias"] = W[p + "b_qkv"]
            sd[h + "self_attention.dense.weight"] = W[p + "w_dense"]
            sd[h + "self_attention.dense.bias"] = W[p + "b_dense"]
            sd[h + "mlp.dense_h_to_4h.weight"] = W[p + "w_up"]
            sd[h + "mlp.dense_h_to_4h.bias"] = W[p + "b_up"]
            sd[h + "mlp.dense_4h_to_h.weight"] = W[p + "w_down"]
            sd[h + "mlp.dense_4h_to_h.bias"] = W[p + "b_down"]
    else:
        sd["gpt_neox.embed_in.weight"] = W["wte"]
        sd["gpt_neox.final_layer_norm.weight"] = W["lnf_g"]
        sd["gpt_neox.final_layer_norm.bias"] = W["lnf_b"]
        sd["embed_out.weight"] = W["lm_head"]
        for l in range(shape.n_layer):
            p, h = f"l{l}.", f"gpt_neox.layers.{l}."
            sd[h + "input_layernorm.weight"] = W[p + "ln1_g"]
            sd[h + "input_layernorm.bias"] = W[p + "ln1_b"]
            sd[h + "post_attention_layernorm.weight"] = W[p + "ln2_g"]
            sd[h + "post_attention_layernorm.bias"] = W[p + "ln2_b"]
            sd[h + "attention.query_key_value.weight"] = W[p + "w_qkv"]
            sd[h + "attention.query_key_value.bias"] = W[p + "b_qkv"]
            sd[h + "attention.dense.weight"] = W[p + "w_dense"]
            sd[h + "attention.dense.bias"] = W[p + "b_dense"]
            sd[h + "mlp.dense_h_to_4h.weight"] = W[p + "w_up"]
            sd[h + "mlp.dense_h_to_4h.bias"] = W[p + "b_up"]
            sd[h + "mlp.dense_4h_to_h.weight"] = W[p + "w_down"]
            sd[h + "mlp.dense_4h_to_h.bias"] = W[p + "b_down"]
    return sd


def build_hf_model(shape: DecoderShape, W: Dict[str, torch.Tensor]):
    """Instantiate the HF model on the meta device and bind ``W`` (fp32, CPU) to it; avoids the
    minutes-long HF random init for 7B shapes."""
    import torch.nn as nn
    from transformers import (BloomForCausalLM, FalconForCausalLM, GPTNeoXForCausalLM, LlamaForCausalLM,
                              OPTForCausalLM)
    cfg = hf_config(shape)
    cls = {ARCH_FALCON: FalconForCausalLM, ARCH_GPT_NEOX: GPTNeoXForCausalLM, ARCH_LLAMA: LlamaForCausalLM,
           ARCH_OPT: OPTForCausalLM, ARCH_BLOOM: BloomForCausalLM}[shape.arch]
    with torch.device("meta"):
        model = cls(cfg)
    sd = {k: v.to(torch.float32).contiguous() for k, v in to_hf_state_dict(shape, W).items()}
    missing, unexpected = model.load_state_dict(sd, strict=False, assign=True)
    # rotary inv_freq buffers are non-persistent and were created on meta: rebuild them
    for mod in model.modules():
        if hasattr(mod, "inv_freq") and getattr(mod.inv_freq, "is_meta", False):
            rope_init = getattr(mod, "rope_init_fn", None)
            if rope_init is None:
                rope_init = type(mod).compute_default_rope_parameters
            inv, _ = rope_init(mod.config, torch.device("cpu"))
            mod.register_buffer("inv_freq", inv, persistent=False)
            if hasattr(mod, "original_inv_freq"):
                mod.register_buffer("original_inv_freq", inv.clone(), persistent=False)
    bad = [m for m in missing if "inv_freq" not in m]
    assert not bad, f"missing keys {bad}"
    assert not unexpected, f"unexpected keys {unexpected}"
    still_meta = [n for n, p in list(model.named_parameters()) + list(model.named_buffers()) if p.is_meta]
    assert not still_meta, f"meta tensors left: {still_meta[:4]}"
    model.eval()
    return model
