"""Plain torch-fp32 restatement of the encoder-decoder forward the reference runs through HF ``transformers``
(test infrastructure only; never imported by the product package).

Follows HF:models/t5/modeling_t5.py:
  T5LayerNorm :46-69 (x * rsqrt(mean(x^2) + eps) * w, no mean subtraction, no bias)
  T5DenseActDense :84-103 (wo(relu(wi(x)))), T5DenseGatedActDense :106-132 (wo(gelu_new(wi_0(x)) * wi_1(x)))
  T5Attention :153-345 (no 1/sqrt(d) scaling; scores += position_bias; fp32 softmax; bucketed relative bias :189-251,
    bidirectional in the encoder, one-sided in the decoder, none in cross-attention :310-317)
  T5Block :411-498 (self-attention, cross-attention, feed-forward, each as x + f(norm(x))), T5Stack :617-790 (the bias of
    layer 0 shared by every layer of the stack, final norm), T5ForConditionalGeneration.forward :1040-1137 (hidden *
    d_model^-0.5 before the head only when the embeddings are tied)
and the reference's procedure around it (analysis/compare_instruct_models.py:178-226): generate from
decoder_start_token_id, one score row per generated position, stop after EOS, top-k scan of the first 10 rows.
Validated against the HF module itself in tests/test_t5_cpu.py; the GPU parity tests call it because it can emulate
the operand rounding of the 16-bit modes.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional

import torch
import torch.nn.functional as F

from llm_interpretation_replication_b200.t5 import EncDecShape, relative_position_bucket
from .decoder_ref import scan_positions


def _norm(x, g, eps):
    return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * g


def _gelu_new(x):
    return 0.5 * x * (1.0 + torch.tanh(0.7978845608028654 * (x + 0.044715 * torch.pow(x, 3.0))))


def _bias(shape: EncDecShape, table: torch.Tensor, q_len: int, k_len: int, bidirectional: bool, offset: int = 0):
    """[nh, q_len, k_len] position bias (HF t5:236-251); queries sit at positions offset .. offset + q_len - 1."""
    ctx = torch.arange(q_len, dtype=torch.long)[:, None] + offset
    mem = torch.arange(k_len, dtype=torch.long)[None, :]
    b = relative_position_bucket(mem - ctx, bidirectional, shape.rel_buckets, shape.rel_max_distance)
    return table[b].permute(2, 0, 1)


def _attend(shape, q, k, v, bias, mask=None):
    """q [Lq, inner], k/v [Lk, inner] -> [Lq, inner]"""
    nh, hd = shape.n_head, shape.head_dim
    qh, kh, vh = q.view(-1, nh, hd), k.view(-1, nh, hd), v.view(-1, nh, hd)
    s = torch.einsum("qhd,khd->hqk", qh, kh)
    if bias is not None:
        s = s + bias
    if mask is not None:
        s = s.masked_fill(~mask[None], float("-inf"))
    a = torch.softmax(s, dim=-1)
    return torch.einsum("hqk,khd->qhd", a, vh).reshape(-1, nh * hd)


def _ff(shape, W, p, x, r, rf=None):
    """``rf``: rounding of the activation that feeds the second projection when it differs from ``r`` (FP16_FFN_SPLIT)."""
    rf = rf if rf is not None else r
    up = F.linear(r(x), W[p + "w_up"])
    if shape.gated:     # rows [0, ff) = wi_0 (gate), [ff, 2ff) = wi_1; formed on the fp32 accumulators
        h = _gelu_new(up[..., :shape.d_ff]) * up[..., shape.d_ff:]
    else:
        h = F.relu(up)
    return F.linear(rf(h), W[p + "w_down"])


def encode(shape: EncDecShape, W: Dict[str, torch.Tensor], ids: torch.Tensor, round_act: Optional[Callable] = None,
           round_attn: Optional[Callable] = None, round_ffn: Optional[Callable] = None) -> torch.Tensor:
    """ids [L] -> encoder output after the final norm, [L, d].  ``round_act`` is applied to every GEMM input,
    ``round_attn`` (default: the same) to q/k/v, as in decoder_ref.forward_hidden."""
    r = round_act if round_act is not None else (lambda t: t)
    ra = round_attn if round_attn is not None else r
    inner, L = shape.inner, ids.shape[0]
    x = W["wte"].float()[ids]
    bias = _bias(shape, W["enc_rel_bias"], L, L, True)
    for l in range(shape.n_enc_layer):
        p = f"e{l}."
        qkv = ra(F.linear(r(_norm(x, W[p + "ln1_g"], shape.norm_eps)), W[p + "w_qkv"]))
        o = _attend(shape, qkv[:, :inner], qkv[:, inner:2 * inner], qkv[:, 2 * inner:], bias)
        x = x + F.linear(r(o), W[p + "w_o"])
        x = x + _ff(shape, W, p, _norm(x, W[p + "ln2_g"], shape.norm_eps), r, round_ffn)
    return _norm(x, W["enc_lnf_g"], shape.norm_eps)


def decode_logits(shape: EncDecShape, W, enc_out: torch.Tensor, dec_ids: torch.Tensor, round_act=None, round_attn=None, round_ffn=None):
    """Logits of the last decoder position given the encoder output and the decoder input ids (start id + generated
    tokens), recomputed without a cache."""
    r = round_act if round_act is not None else (lambda t: t)
    ra = round_attn if round_attn is not None else r
    inner, Ld = shape.inner, dec_ids.shape[0]
    x = W["wte"].float()[dec_ids]
    causal = torch.tril(torch.ones(Ld, Ld, dtype=torch.bool))
    bias = _bias(shape, W["dec_rel_bias"], Ld, Ld, False)
    enc_r = r(enc_out)
    for l in range(shape.n_dec_layer):
        p = f"d{l}."
        qkv = ra(F.linear(r(_norm(x, W[p + "ln1_g"], shape.norm_eps)), W[p + "w_qkv"]))
        o = _attend(shape, qkv[:, :inner], qkv[:, inner:2 * inner], qkv[:, 2 * inner:], bias, causal)
        x = x + F.linear(r(o), W[p + "w_o"])
        cq = ra(F.linear(r(_norm(x, W[p + "ln2_g"], shape.norm_eps)), W[p + "w_cq"]))
        ckv = ra(F.linear(enc_r, W[p + "w_ckv"]))
        o = _attend(shape, cq, ckv[:, :inner], ckv[:, inner:], None)
        x = x + F.linear(r(o), W[p + "w_co"])
        x = x + _ff(shape, W, p, _norm(x, W[p + "ln3_g"], shape.norm_eps), r, round_ffn)
    h = _norm(x[-1], W["dec_lnf_g"], shape.norm_eps)
    if shape.tied_head:
        h = h * (shape.d_model ** -0.5)
    return F.linear(r(h), W["lm_head"].float()).float()


def greedy_scores(shape, W, ids: List[int], n_positions: int, eos_id: int, round_act=None, round_attn=None, round_ffn=None):
    """scores[t] = logits after the start id + t generated tokens (HF generate, greedy); stops after the position whose
    argmax is EOS, as ``generate`` does."""
    enc = encode(shape, W, torch.tensor(ids, dtype=torch.long), round_act, round_attn, round_ffn)
    dec = [shape.decoder_start_id]
    rows, gen = [], []
    for _ in range(n_positions):
        row = decode_logits(shape, W, enc, torch.tensor(dec, dtype=torch.long), round_act, round_attn, round_ffn)
        rows.append(row)
        nxt = int(torch.argmax(row))
        gen.append(nxt)
        dec.append(nxt)
        if nxt == eos_id:
            break
    return rows, gen


def score_ids(shape, W, ids: List[int], id_a: int, id_b: int, topk: int = 2, max_lookahead: int = 10,
              eos_id: Optional[int] = None, round_act=None, round_attn=None, round_ffn=None):
    eos = shape.eos_id if eos_id is None else eos_id
    rows, gen = greedy_scores(shape, W, ids, max_lookahead, eos, round_act, round_attn, round_ffn)
    pa, pb, pos, found = scan_positions(rows, id_a, id_b, topk)
    row = rows[pos]
    return dict(p_a=pa, p_b=pb, rel=(pa / (pa + pb) if pa + pb > 0 else float("nan")), lse=torch.logsumexp(row, dim=-1).item(),
                logit_a=row[id_a].item(), logit_b=row[id_b].item(), pos_found=pos, found=found, gen_ids=gen, n_scores=len(rows))
