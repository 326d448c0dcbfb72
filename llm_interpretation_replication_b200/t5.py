"""Encoder-decoder (T5 family) side of the scoring path: shapes, the flat weight layout and the host-side scorer.

The reference reaches these models through ``AutoModelForSeq2SeqLM`` and the ``is_encoder_decoder`` branch of
``get_yes_no_logprobs`` (reference analysis/compare_instruct_models.py:178-226,470-476 for ``allenai/tk-instruct-3b-def``
and ``bigscience/T0_3B``; compare_base_vs_instruct.py:137-138,192-237 and run_base_vs_instruct_100q.py:290-325,447-448 for
``google/t5-v1_1-base`` / ``google/flan-t5-base``).  :class:`EncDecScorer` keeps the :class:`~.scorer.Scorer` interface
(``score_ids``, ``score_prompts`` work unchanged) over the ``b2s_encdec_*`` entry points of include/b200score.h.

Slot order (``B2S_T5_*`` in the header):
  global:      0 wte[V,d]  1 enc_lnf_g[d]  2 dec_lnf_g[d]  3 lm_head[V,d]  4 enc_bias[nh, 2*max_pos-1]  5 dec_bias[...]  (6, 7 reserved)
  per encoder layer (8):  ln1_g, w_qkv[3*inner,d] (q rows, k rows, v rows), w_o[d,inner], ln2_g, w_up, w_down[d,ff]
  per decoder layer (12): ln1_g, w_qkv, w_o, ln2_g, w_cq[inner,d], w_ckv[2*inner,d] (k rows, v rows), w_co[d,inner], ln3_g,
                          w_up, w_down
  w_up = wi[ff,d] (ReLU models) or [wi_0 rows; wi_1 rows] = [2*ff,d] (gated tanh-GELU: T5 v1.1, Flan-T5, LM-adapted T5).
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch

from . import _lib
from .shapes import MODE_BF16_FAST, MODE_BF16_SPLIT, MODE_FP16, MODE_FP16_FFN_SPLIT, MODE_FP32_CHECK
from .scorer import Scorer, MAX_LOOK_AHEAD, fit_batch

FF_RELU, FF_GATED_GELU, FF_GATED_GELU_FUSED = 0, 1, 2   # must match B2S_T5_FF_* in include/b200score.h
N_GLOBAL, N_PER_ENC, N_PER_DEC = 8, 8, 12


@dataclass(frozen=True)
class EncDecShape:
    name: str
    n_enc_layer: int
    n_dec_layer: int
    d_model: int
    n_head: int
    d_ff: int
    vocab: int
    head_dim: int = 64
    norm_eps: float = 1e-6
    gated: bool = True            # feed_forward_proj "gated-gelu" (v1.1 / Flan / LM-adapted) vs "relu" (original T5)
    rel_buckets: int = 32
    rel_max_distance: int = 128
    tied_head: bool = False       # original T5: lm_head is the embedding and the decoder output is scaled by d^-0.5 first
                                  # (HF t5:1107-1108 `scale_decoder_outputs`); v1.1 / Flan / LM-adapted: neither
    eos_id: int = 1
    pad_id: int = 0
    decoder_start_id: int = 0
    max_pos: int = 512            # longest prompt the bias tables cover (tokenizer model_max_length of the T5 family)
    is_encoder_decoder: bool = True

    @property
    def inner(self) -> int:
        return self.n_head * self.head_dim

    @property
    def up_dim(self) -> int:
        return self.d_ff * (2 if self.gated else 1)


def t5_v1_1_base() -> EncDecShape:
    """google/t5-v1_1-base and google/flan-t5-base (public config.json: d_model 768, d_kv 64, d_ff 2048, 12 + 12 layers,
    12 heads, vocab 32128, gated-gelu, untied head)."""
    return EncDecShape("t5-v1_1-base", 12, 12, 768, 12, 2048, 32128)


def t5_xl_lm_adapt() -> EncDecShape:
    """bigscience/T0_3B and allenai/tk-instruct-3b-def: both fine-tuned from google/t5-xl-lm-adapt (d_model 2048, d_kv 64,
    d_ff 5120, 24 + 24 layers, 32 heads, gated-gelu, untied head).  Vocabulary 32128 as T5 v1.1 ships it; a checkpoint whose
    config.json says otherwise (resized embeddings) is taken at its word by checkpoint.load_t5."""
    return EncDecShape("t5-xl-lm-adapt", 24, 24, 2048, 32, 5120, 32128)


def tiny_t5(vocab=296, n_layer=2, n_head=3, gated=True, d_ff=256) -> EncDecShape:
    return EncDecShape("tiny-t5" if gated else "tiny-t5-relu", n_layer, n_layer, 64 * n_head, n_head, d_ff, vocab, gated=gated,
                       max_pos=256)


def tiny_t5_relu() -> EncDecShape:
    """Original-T5 feed-forward (ReLU, no gate) with more heads than d_model / 64 would give (inner != d_model)."""
    return EncDecShape("tiny-t5-relu", 3, 3, 128, 4, 192, 280, gated=False, max_pos=256)


REGISTRY = {"t5-v1_1-base": t5_v1_1_base, "flan-t5-base": t5_v1_1_base, "t5-xl-lm-adapt": t5_xl_lm_adapt,
            "tiny-t5": tiny_t5, "tiny-t5-relu": tiny_t5_relu}
HUB_TO_SHAPE = {"google/t5-v1_1-base": t5_v1_1_base, "google/flan-t5-base": t5_v1_1_base,
                "bigscience/T0_3B": t5_xl_lm_adapt, "allenai/tk-instruct-3b-def": t5_xl_lm_adapt}


# ---- relative position bias ---------------------------------------------------------------------
def relative_position_bucket(relative_position: torch.Tensor, bidirectional: bool, num_buckets: int, max_distance: int) -> torch.Tensor:
    """Bucket of (key position - query position), the same torch operations in the same order as
    HF:models/t5/modeling_t5.py:189-234 (T5Attention._relative_position_bucket), so that the float32 logarithm lands on
    the same side of every bucket boundary as the model the reference loads."""
    relative_buckets = torch.zeros_like(relative_position)
    if bidirectional:
        num_buckets //= 2
        relative_buckets = relative_buckets + (relative_position > 0).to(torch.long) * num_buckets
        relative_position = torch.abs(relative_position)
    else:
        relative_position = -torch.min(relative_position, torch.zeros_like(relative_position))
    max_exact = num_buckets // 2
    is_small = relative_position < max_exact
    if_large = max_exact + (torch.log(relative_position.float() / max_exact) / math.log(max_distance / max_exact)
                            * (num_buckets - max_exact)).to(torch.long)
    if_large = torch.min(if_large, torch.full_like(if_large, num_buckets - 1))
    return relative_buckets + torch.where(is_small, relative_position, if_large)


def bias_table(rel_bias: torch.Tensor, bidirectional: bool, shape: EncDecShape) -> torch.Tensor:
    """relative_attention_bias.weight [buckets, n_head] -> fp32 [n_head, 2*max_pos - 1], entry [h][(key - query) + max_pos - 1]
    (HF t5:236-251 compute_bias, laid out by distance instead of by (query, key))."""
    R = shape.max_pos - 1
    rel = torch.arange(-R, R + 1, dtype=torch.long)
    b = relative_position_bucket(rel, bidirectional, shape.rel_buckets, shape.rel_max_distance)
    return rel_bias.detach().float().cpu()[b].t().contiguous()


# ---- weights ------------------------------------------------------------------------------------
ENC_SLOTS = ["ln1_g", "w_qkv", "w_o", "ln2_g", "w_up", "w_down", None, None]
DEC_SLOTS = ["ln1_g", "w_qkv", "w_o", "ln2_g", "w_cq", "w_ckv", "w_co", "ln3_g", "w_up", "w_down", None, None]


def slot_names(shape: EncDecShape) -> List[Optional[str]]:
    names: List[Optional[str]] = ["wte", "enc_lnf_g", "dec_lnf_g", "lm_head", "enc_bias", "dec_bias", None, None]
    for l in range(shape.n_enc_layer):
        names += [None if s is None else f"e{l}.{s}" for s in ENC_SLOTS]
    for l in range(shape.n_dec_layer):
        names += [None if s is None else f"d{l}.{s}" for s in DEC_SLOTS]
    return names


def is_matrix(name: str) -> bool:
    return name.split(".")[-1] in ("wte", "lm_head", "w_qkv", "w_o", "w_cq", "w_ckv", "w_co", "w_up", "w_down")


def init_weights(shape: EncDecShape, seed: int = 0, device: str = "cpu", bf16_representable: bool = True,
                 weight_dtype=None) -> Dict[str, torch.Tensor]:
    """Random weights at the scales of HF's T5 ``_init_weights`` (HF t5:529-600 with factor 1: q ~ N(0, (d*d_kv)^-1/2),
    k, v, wi ~ N(0, d^-1/2), o ~ N(0, inner^-1/2), wo ~ N(0, ff^-1/2)), so that the unscaled q.k scores, the bias and the
    gate all carry signal; the embedding is N(0, 1) as HF's, the head N(0, d^-1/2) (unit-variance logits), the norm gains
    and the relative bias random.  With ``weight_dtype`` the matrices are produced directly in that type (full-size
    models on the GPU); otherwise fp32 masters, rounded to bf16-representable values when asked."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    d, ff, V, inner, dk = shape.d_model, shape.d_ff, shape.vocab, shape.inner, shape.head_dim

    def mat(r, c, std):
        w = torch.empty(r, c, dtype=torch.float32, device=device).normal_(0.0, std, generator=g)
        if weight_dtype is not None:
            return w.to(weight_dtype)
        return w.to(torch.bfloat16).to(torch.float32) if bf16_representable else w

    def gain(n):
        return 1.0 + 0.1 * torch.empty(n, dtype=torch.float32, device=device).normal_(0.0, 1.0, generator=g)

    def qkv():
        return torch.cat((mat(inner, d, (d * dk) ** -0.5), mat(inner, d, d ** -0.5), mat(inner, d, d ** -0.5)), dim=0)

    W: Dict[str, torch.Tensor] = {"wte": mat(V, d, 1.0), "enc_lnf_g": gain(d), "dec_lnf_g": gain(d)}
    W["lm_head"] = W["wte"] if shape.tied_head else mat(V, d, d ** -0.5)
    W["enc_rel_bias"] = torch.empty(shape.rel_buckets, shape.n_head, dtype=torch.float32, device=device).normal_(0.0, 0.5, generator=g)
    W["dec_rel_bias"] = torch.empty(shape.rel_buckets, shape.n_head, dtype=torch.float32, device=device).normal_(0.0, 0.5, generator=g)
    for l in range(shape.n_enc_layer):
        p = f"e{l}."
        W[p + "ln1_g"] = gain(d); W[p + "w_qkv"] = qkv(); W[p + "w_o"] = mat(d, inner, inner ** -0.5)
        W[p + "ln2_g"] = gain(d); W[p + "w_up"] = mat(shape.up_dim, d, d ** -0.5); W[p + "w_down"] = mat(d, ff, ff ** -0.5)
    for l in range(shape.n_dec_layer):
        p = f"d{l}."
        W[p + "ln1_g"] = gain(d); W[p + "w_qkv"] = qkv(); W[p + "w_o"] = mat(d, inner, inner ** -0.5)
        W[p + "ln2_g"] = gain(d); W[p + "w_cq"] = mat(inner, d, (d * dk) ** -0.5)
        W[p + "w_ckv"] = torch.cat((mat(inner, d, d ** -0.5), mat(inner, d, d ** -0.5)), dim=0)
        W[p + "w_co"] = mat(d, inner, inner ** -0.5); W[p + "ln3_g"] = gain(d)
        W[p + "w_up"] = mat(shape.up_dim, d, d ** -0.5); W[p + "w_down"] = mat(d, ff, ff ** -0.5)
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


def pointer_table(shape: EncDecShape, Wdev: Dict[str, torch.Tensor]):
    names = slot_names(shape)
    arr = (C.c_void_p * len(names))()
    for i, n in enumerate(names):
        t = Wdev.get(n) if n is not None else None
        arr[i] = None if t is None else t.data_ptr()
    return arr


# ---- host-side scorer ---------------------------------------------------------------------------
class EncDecScorer(Scorer):
    """One T5-family model resident on one GPU, behind the interface of :class:`~.scorer.Scorer`: ``score_ids`` (and with
    it ``score_prompts``, ``dist.score_ids_sharded``) runs the encoder-decoder entry points instead.  ``weights_dev`` is
    the flat dict of :func:`init_weights` / ``checkpoint.load_t5`` on the device (matrices in the mode's weight type).

    Default mode: FP16_FFN_SPLIT.  The T5 v1.1 family is the textbook case of activations that leave fp16's range, and the place
    is known: the product act(wi_0 x) * wi_1 x that feeds ``wo`` (HF keeps ``wo`` in fp32 for fp16 inference for that reason,
    HF t5:527 ``_keep_in_fp32_modules``).  The default therefore runs fp16 everywhere except there: that activation is stored as
    bf16 hi + lo halves (formed on the fp32 accumulators) and ``wo`` stays bf16; its output lands in the fp32 residual stream
    like every projection's.  One GEMM in four pays the doubled K.  BF16_SPLIT (every GEMM split, bf16 weights untouched) and
    plain FP16 (an overflow is flagged as B2S_ERANGE) remain available.

    Batch size: the one-token decoder steps are launch- and latency-bound below ~1000 rows, so the default batch is 131072
    prompt tokens (about 1200 prompts; 37 GB of workspace at T0_3B size, dominated by the resident cross-attention K/V:
    tools/t5_bench.py measured 1704 / 2284 / 2341 prompts/s at 32768 / 131072 / 262144 tokens)."""

    is_encoder_decoder = True

    def __init__(self, shape: EncDecShape, weights_dev: Dict[str, torch.Tensor], mode: int = MODE_FP16_FFN_SPLIT,
                 device: Optional[torch.device] = None, max_tokens: Optional[int] = 131072, max_seqs: int = 2048,
                 max_lookahead: int = MAX_LOOK_AHEAD):
        if not torch.cuda.is_available():
            raise _lib.B2SError("no CUDA device: the B200 scorer has no CPU fallback")
        self.lib = _lib.load()
        self.shape = shape
        self.mode = mode
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.max_tokens, self.max_seqs, self.max_lookahead = max_tokens, int(max_seqs), int(max_lookahead)
        want = {MODE_FP32_CHECK: torch.float32, MODE_FP16: torch.float16, MODE_FP16_FFN_SPLIT: torch.float16}.get(mode, torch.bfloat16)
        W = dict(weights_dev)
        if mode in (MODE_FP16, MODE_FP16_FFN_SPLIT):
            # FP16_FFN_SPLIT: the second feed-forward projection stays bf16 (it meets the bf16 hi + lo halves of the activation)
            keep = (lambda k: k.endswith(".w_down")) if mode == MODE_FP16_FFN_SPLIT else (lambda k: False)
            W = {k: (v.to(torch.bfloat16 if keep(k) else torch.float16) if is_matrix(k) else v) for k, v in W.items()}
            if weights_dev.get("lm_head") is weights_dev.get("wte"):
                W["lm_head"] = W["wte"]
        ff_type = FF_GATED_GELU if shape.gated else FF_RELU
        if shape.gated and mode != MODE_FP32_CHECK:
            # tensor-core modes: wi_0 / wi_1 rows interleaved, gelu(wi_0 x) * wi_1 x formed in the projection's epilogue
            ff_type = FF_GATED_GELU_FUSED
            ff = shape.d_ff
            for k in [k for k in W if k.endswith(".w_up")]:
                w = W[k]
                W[k] = torch.stack((w[:ff], w[ff:]), dim=1).reshape(2 * ff, w.shape[1]).contiguous()
        if shape.tied_head:   # logits = lm_head(hidden * d^-0.5) (HF t5:1105-1108): folded into the final gain
            W["dec_lnf_g"] = W["dec_lnf_g"] * (shape.d_model ** -0.5)
        W["enc_bias"] = bias_table(W.pop("enc_rel_bias"), True, shape).to(self.device)
        W["dec_bias"] = bias_table(W.pop("dec_rel_bias"), False, shape).to(self.device)
        for k, v in W.items():
            assert v.is_cuda and v.is_contiguous(), k
            exp = torch.bfloat16 if (mode == MODE_FP16_FFN_SPLIT and k.endswith(".w_down")) else want
            assert v.dtype == (exp if is_matrix(k) else torch.float32), f"{k} is {v.dtype}"
        self.W = W
        self._ptrs = pointer_table(shape, W)
        d = _lib.EncDecDesc(abi_version=_lib.ABI_VERSION, n_enc_layer=shape.n_enc_layer, n_dec_layer=shape.n_dec_layer,
                            d_model=shape.d_model, n_head=shape.n_head, head_dim=shape.head_dim, d_ff=shape.d_ff,
                            vocab=shape.vocab, norm_eps=shape.norm_eps, ff_type=ff_type, max_pos=shape.max_pos,
                            compute_mode=mode, decoder_start_id=shape.decoder_start_id)
        self._handle = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2s_encdec_create(C.byref(d), self._ptrs, len(self._ptrs), self.device.index or 0,
                                                  C.byref(self._handle)), "b2s_encdec_create")
        if self.max_tokens is None:
            torch.cuda.empty_cache()
            free, _ = torch.cuda.mem_get_info(self.device)
            self.max_tokens, self.max_seqs = fit_batch(lambda t, s, g: self.lib.b2s_encdec_workspace_bytes(self._handle, t, s, g), free,
                                                       self.max_lookahead, cap_tokens=131072)
        self.max_tokens = int(self.max_tokens)
        nbytes = self.lib.b2s_encdec_workspace_bytes(self._handle, self.max_tokens, self.max_seqs, self.max_lookahead)
        if nbytes == 0:
            raise _lib.B2SError("b2s_encdec_workspace_bytes returned 0")
        self.workspace = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        self.workspace_bytes = nbytes

    def close(self):
        if getattr(self, "_handle", None) is not None and self._handle:
            self.lib.b2s_encdec_destroy(self._handle)
            self._handle = None
        self.workspace = None
        self.W = None
        self._ptrs = None

    def score_batch_host(self, t_tok, t_cu, t_cand, topk, max_lookahead, eos_id, out, gen, gen_tokens=None):
        n = t_cu.numel() - 1
        G = max_lookahead if gen_tokens is None else gen_tokens
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2s_encdec_score_host(self._handle, _lib.ptr(t_tok), _lib.ptr(t_cu), n, _lib.ptr(t_cand), topk,
                                                      max_lookahead, G, eos_id, _lib.ptr(out), _lib.ptr(gen),
                                                      _lib.ptr(self.workspace), self.workspace_bytes, _lib.current_stream()),
                       "b2s_encdec_score_host")

    def score_batch_device(self, d_tok, d_cu, d_cand, n, total, max_len, topk, max_lookahead, eos_id, d_out, d_gen):
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2s_encdec_score(self._handle, _lib.ptr(d_tok), _lib.ptr(d_cu), n, total, max_len, _lib.ptr(d_cand),
                                                 topk, max_lookahead, max_lookahead, eos_id, _lib.ptr(d_out), _lib.ptr(d_gen),
                                                 _lib.ptr(self.workspace), self.workspace_bytes, _lib.current_stream()),
                       "b2s_encdec_score")
