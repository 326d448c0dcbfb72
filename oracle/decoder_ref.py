"""Plain torch-fp32 restatement of the decoder forward the reference runs through HF
``transformers`` (test infrastructure only; never imported by the product package).

Follows, for the two families on the hot path:
  Falcon   HF:models/falcon/modeling_falcon.py:60-99 (linear, rotate_half RoPE), :154-165 (cos/sin),
           :285-288 (MQA qkv split), :316-378 (attention), :531-544 (MLP, exact GELU),
           :580-636 (parallel-attn layer, single input_layernorm), :700-834 (model), :905-906 (head)
  GPT-NeoX HF:models/gpt_neox/modeling_gpt_neox.py:38-49 (MLP), :126-159 (partial rotary),
           :203-244 (attention, per-head [q,k,v] interleave), :258-289 (layer, parallel/sequential),
           :331-381 (model), :463-464 (embed_out)
  Qwen     the Llama block with a bias on q|k|v (HF:models/qwen2/modeling_qwen2.py Qwen2Attention q/k/v_proj bias=True; the
           first-generation Qwen's c_attn bias)
  Llama    HF:models/llama/modeling_llama.py LlamaRMSNorm (x * rsqrt(mean(x^2) + eps) * w), LlamaMLP
           (down(silu(gate(x)) * up(x))), LlamaAttention (grouped-query heads, rotate_half RoPE on the whole head),
           LlamaDecoderLayer (sequential residual, two norms) -- also Mistral without a sliding window
  OPT      HF:models/opt/modeling_opt.py OPTLearnedPositionalEmbedding (positions + 2), OPTDecoderLayer with
           do_layer_norm_before (pre-LN, sequential residual), ReLU MLP, q/k/v/out projections with biases, no rotary
  BLOOM    HF:models/bloom/modeling_bloom.py word_embeddings_layernorm, build_alibi_tensor (slope[h] * key position
           added to the scaled scores), per-head [q,k,v] fused projection, bloom_gelu_forward (tanh GELU), BloomBlock
It is validated against the HF modules themselves in tests/test_oracle_cpu.py (test_torch_decoder_restatement_matches_reference), and is what the
GPU parity tests call (it is faster than ``generate`` and can emulate bf16 operand rounding).
"""
from __future__ import annotations

import math
from typing import Callable, Dict, List, Optional

import torch
import torch.nn.functional as F

from llm_interpretation_replication_b200.shapes import (DecoderShape, QKV_FALCON_MQA, NORM_RMS, MLP_SWIGLU, POS_ROPE, POS_LEARNED,
                                                        POS_ALIBI, ACT_RELU, ACT_GELU_TANH)


def _rope_cos_sin(shape: DecoderShape, L: int):
    rd = shape.rotary_dim
    inv_freq = 1.0 / (shape.rope_theta ** (torch.arange(0, rd, 2, dtype=torch.int64).float() / rd))
    pos = torch.arange(L, dtype=torch.float32)
    freqs = pos[:, None] * inv_freq[None, :]
    emb = torch.cat((freqs, freqs), dim=-1)
    return emb.cos(), emb.sin()          # [L, rd]


def _rotate_half(x):
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


def _apply_rope(x, cos, sin, rd):
    # x [L, H, hd]; rotate the first rd dims
    xr, xp = x[..., :rd], x[..., rd:]
    xr = xr * cos[:, None, :] + _rotate_half(xr) * sin[:, None, :]
    return torch.cat((xr, xp), dim=-1)


def _split_qkv(shape: DecoderShape, qkv):
    L = qkv.shape[0]
    nh, nkv, hd = shape.n_head, shape.n_kv_head, shape.head_dim
    if shape.qkv_layout == QKV_FALCON_MQA:
        x = qkv.view(L, nh + 2 * nkv, hd)
        return x[:, :nh], x[:, nh:nh + nkv], x[:, nh + nkv:nh + 2 * nkv]
    x = qkv.view(L, nh, 3, hd)
    return x[:, :, 0], x[:, :, 1], x[:, :, 2]


def forward_hidden(shape: DecoderShape, W: Dict[str, torch.Tensor], ids: torch.Tensor,
                   round_act: Optional[Callable] = None, dtype=torch.float32, round_attn: Optional[Callable] = None,
                   linear_fn: Optional[Callable] = None) -> torch.Tensor:
    """ids [L] -> hidden after the final LayerNorm, [L, d].  ``round_act`` (e.g. bf16 round trip) is
    applied to every GEMM input and to q/k/v, emulating bf16 MMA operands with fp32 accumulate; ``round_attn``
    overrides it for q/k/v (the BF16_SPLIT mode keeps those in fp16 while the GEMM inputs are hi + lo bf16 halves).
    ``linear_fn(x, weight, bias)`` replaces the four linear layers of every block (oracle/int8_ref.py: bitsandbytes'
    Linear8bitLt restated); the LM head is not one of them."""
    r = round_act if round_act is not None else (lambda t: t)
    lin = linear_fn if linear_fn is not None else (lambda t, w, b: F.linear(r(t), w, b))
    ra = round_attn if round_attn is not None else r
    d, nh, nkv, hd = shape.d_model, shape.n_head, shape.n_kv_head, shape.head_dim
    L = ids.shape[0]
    rope = shape.pos_type == POS_ROPE
    if rope:
        cos, sin = _rope_cos_sin(shape, L)
        cos, sin = cos.to(dtype), sin.to(dtype)
    x = W["wte"].to(dtype)[ids]
    if shape.pos_type == POS_LEARNED:
        x = x + W["wpe"].to(dtype)[torch.arange(L) + shape.pos_offset]
    if shape.emb_norm:
        x = F.layer_norm(x, (d,), W["emb_ln_g"].to(dtype), W["emb_ln_b"].to(dtype), shape.norm_eps)
    causal = torch.tril(torch.ones(L, L, dtype=torch.bool))
    alibi = None
    if shape.pos_type == POS_ALIBI:      # [nh, 1, L]: slope * key position, the same for every query row
        alibi = (W["alibi_slopes"].to(dtype)[:, None] * torch.arange(L, dtype=dtype)[None, :])[:, None, :]

    def norm(t, g, b):
        if shape.norm_type == NORM_RMS:
            return t * torch.rsqrt(t.pow(2).mean(-1, keepdim=True) + shape.norm_eps) * W[g].to(dtype)
        return F.layer_norm(t, (d,), W[g].to(dtype), W[b].to(dtype), shape.norm_eps)

    def mlp_act(t, p, bias):
        up = lin(t, W[p + "w_up"].to(dtype), bias("b_up"))
        if shape.mlp_type == MLP_SWIGLU:        # rows [0, ff) = gate_proj, [ff, 2ff) = up_proj; formed on the fp32 accumulators
            return F.silu(up[..., :shape.d_ff]) * up[..., shape.d_ff:]
        if shape.act_type == ACT_RELU:
            return F.relu(up)
        if shape.act_type == ACT_GELU_TANH:     # applied to the fp32 accumulators in the up-projection's epilogue
            return up * 0.5 * (1.0 + torch.tanh(0.79788456 * up * (1 + 0.044715 * up * up)))
        return F.gelu(up)

    for l in range(shape.n_layer):
        p = f"l{l}."
        bias = (lambda n: W[p + n].to(dtype)) if shape.has_bias else \
               (lambda n: W[p + n].to(dtype) if (n == "b_qkv" and shape.qkv_bias) else None)
        ln1 = norm(x, p + "ln1_g", p + "ln1_b")
        qkv = lin(ln1, W[p + "w_qkv"].to(dtype), bias("b_qkv"))
        q, k, v = _split_qkv(shape, qkv)
        if rope:
            q, k = _apply_rope(q, cos, sin, shape.rotary_dim), _apply_rope(k, cos, sin, shape.rotary_dim)
        q, k = ra(q), ra(k)
        v = ra(v)
        rep = nh // k.shape[1]
        kk = k.repeat_interleave(rep, dim=1)
        vv = v.repeat_interleave(rep, dim=1)
        s = torch.einsum("qhd,khd->hqk", q, kk) / math.sqrt(hd)
        if alibi is not None:
            s = s + alibi
        s = s.masked_fill(~causal[None], float("-inf"))
        a = torch.softmax(s, dim=-1)
        o = torch.einsum("hqk,khd->qhd", a, vv).reshape(L, nh * hd)
        attn = lin(o, W[p + "w_dense"].to(dtype), bias("b_dense"))
        if shape.parallel_residual:
            ln2 = ln1 if shape.n_ln_per_layer == 1 else norm(x, p + "ln2_g", p + "ln2_b")
            h = mlp_act(ln2, p, bias)
            mlp = lin(h, W[p + "w_down"].to(dtype), bias("b_down"))
            x = mlp + attn + x
        else:
            x1 = attn + x
            ln2 = norm(x1, p + "ln2_g", p + "ln2_b")
            h = mlp_act(ln2, p, bias)
            mlp = lin(h, W[p + "w_down"].to(dtype), bias("b_down"))
            x = mlp + x1
    return norm(x, "lnf_g", "lnf_b")


def last_logits(shape, W, ids, round_act=None, dtype=torch.float32, round_attn=None, linear_fn=None) -> torch.Tensor:
    r = round_act if round_act is not None else (lambda t: t)
    h = forward_hidden(shape, W, ids, round_act, dtype, round_attn, linear_fn)[-1]
    return F.linear(r(h), W["lm_head"].to(dtype)).float()


def bf16_round(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(t.dtype)


def scan_positions(score_rows: List[torch.Tensor], id_a: int, id_b: int, topk: int):
    """The reference's look-ahead rule (analysis/run_base_vs_instruct_100q.py:348-363;
    compare_instruct_models.py:249-272): first of the (<=10) positions where either candidate id is
    among the top-k probabilities; both probabilities are read there; otherwise position 0."""
    for pos, row in enumerate(score_rows):
        probs = torch.softmax(row, dim=-1)
        _, top = torch.topk(probs, k=topk)
        if id_a in top or id_b in top:
            return probs[id_a].item(), probs[id_b].item(), pos, True
    probs = torch.softmax(score_rows[0], dim=-1)
    return probs[id_a].item(), probs[id_b].item(), 0, False


def greedy_scores(shape, W, ids: List[int], n_positions: int, eos_id: int, round_act=None,
                  dtype=torch.float32, round_attn=None, linear_fn=None):
    """Greedy decode restated without a KV cache: scores[t] = logits after prompt + t generated
    tokens (HF:generation/utils.py:2743-2805; SURVEY.md Appendix A, last row).  Stops after the
    position whose argmax is EOS, as ``generate`` does."""
    seq = list(ids)
    rows, gen = [], []
    for _ in range(n_positions):
        row = last_logits(shape, W, torch.tensor(seq, dtype=torch.long), round_act, dtype, round_attn, linear_fn)
        rows.append(row)
        nxt = int(torch.argmax(row))
        gen.append(nxt)
        seq.append(nxt)
        if nxt == eos_id:
            break
    return rows, gen


def score_ids(shape, W, ids: List[int], id_a: int, id_b: int, topk: int = 5, max_lookahead: int = 10,
              eos_id: Optional[int] = None, round_act=None, dtype=torch.float32, round_attn=None, linear_fn=None):
    eos = shape.eos_id if eos_id is None else eos_id
    rows, gen = greedy_scores(shape, W, ids, max_lookahead, eos, round_act, dtype, round_attn, linear_fn)
    pa, pb, pos, found = scan_positions(rows, id_a, id_b, topk)
    row = rows[pos]
    lse = torch.logsumexp(row, dim=-1).item()
    return dict(p_a=pa, p_b=pb, rel=(pa / (pa + pb) if pa + pb > 0 else float("nan")), lse=lse,
                logit_a=row[id_a].item(), logit_b=row[id_b].item(), pos_found=pos, found=found,
                gen_ids=gen, n_scores=len(rows))
