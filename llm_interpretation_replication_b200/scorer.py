"""Host-side mirror of the reference's scoring interface on top of libb200score.

Kept names / return schemas:
  * ``get_yes_no_logprobs(model, tokenizer, prompt, device)``  -- reference
    analysis/run_base_vs_instruct_100q.py:279-392 (variant "100q"), compare_instruct_models.py:171-293
    (variant "instruct"), compare_base_vs_instruct.py:185-305 (variant "base_vs").  Here ``model`` is a
    :class:`Scorer`; the call is a thin wrapper over the batched :func:`score_prompts`.
  * ``score_prompts`` replaces the reference's ``for prompt in prompts`` loop
    (run_base_vs_instruct_100q.py:464-476) with length-bucketed, varlen-packed batches.
PyTorch only owns memory here (weights, workspace, pinned staging); all arithmetic happens inside the
C-ABI library.  There is no CPU path: a missing library or GPU raises.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from .shapes import DecoderShape, MODE_BF16_FAST, MODE_FP16, MODE_FP32_CHECK, MODE_INT8_EMU, MLP_SWIGLU, MLP_SWIGLU_FUSED
from . import weights as _weights

MAX_LOOK_AHEAD = 10   # reference run_base_vs_instruct_100q.py:285

VARIANTS = {
    # variant: (topk, candidate ids resolved with special tokens?, relative_prob fallback)
    "100q": dict(topk=5, cand_specials=False, rel_fallback=0.5),
    "instruct": dict(topk=2, cand_specials=True, rel_fallback=float("nan")),
    "base_vs": dict(topk=2, cand_specials=True, rel_fallback=None),
}


def candidate_ids(tokenizer, variant: str, yes: str = " Yes", no: str = " No") -> Tuple[int, int]:
    """Candidate-token ids exactly as each reference entry point resolves them:
    run_base_vs_instruct_100q.py:332-335 (no special tokens) vs compare_instruct_models.py:229-233
    (special tokens included, so BOS-prepending tokenizers yield yes == no == BOS)."""
    if VARIANTS[variant]["cand_specials"]:
        y, n = tokenizer(yes).input_ids, tokenizer(no).input_ids
    else:
        y = tokenizer(yes, add_special_tokens=False).input_ids
        n = tokenizer(no, add_special_tokens=False).input_ids
    y_id = y[0] if len(y) > 0 else tokenizer.encode(yes.strip())[0]
    n_id = n[0] if len(n) > 0 else tokenizer.encode(no.strip())[0]
    return int(y_id), int(n_id)


def int8_rowwise_dequant(w: torch.Tensor) -> torch.Tensor:
    """fp16(rint(W * 127 / s) * s / 127) with s = max |W[row]|: the weight a bitsandbytes ``Linear8bitLt`` effectively multiplies
    by (``CB`` int8 rows, ``SCB`` row scales; ``F.double_quant`` / ``int8_vectorwise_quant`` with round-to-nearest-even)."""
    wf = w.float()
    s = wf.abs().amax(dim=1, keepdim=True).clamp_min(1e-30)
    return (torch.round(wf * (torch.full_like(s, 127.0) / s)) * (s / 127.0)).to(torch.float16)


def plan_batches(lengths: Sequence[int], max_tokens: int, max_seqs: int) -> List[List[int]]:
    """Length-bucketed packing: prompts sorted by length (longest first, stable) and cut into consecutive batches of at
    most ``max_tokens`` packed tokens and ``max_seqs`` prompts.  The number of batches is the smallest the greedy cut
    reaches; among those plans the one with the smallest token cap is taken, so the batches carry near-equal token
    counts (a full batch followed by a small remainder would run its one-token steps with a nearly empty GEMM wave).
    A prompt's result does not depend on the batch it lands in."""
    order = sorted(range(len(lengths)), key=lambda i: (-lengths[i], i))
    if not order:
        return []
    longest = lengths[order[0]]
    if longest > max_tokens:
        raise ValueError(f"prompt {order[0]} has {longest} tokens, more than max_tokens={max_tokens}")

    def cut(cap: int) -> List[List[int]]:
        batches, cur, cur_tok = [], [], 0
        for i in order:
            L = lengths[i]
            if cur and (cur_tok + L > cap or len(cur) >= max_seqs):
                batches.append(cur)
                cur, cur_tok = [], 0
            cur.append(i)
            cur_tok += L
        batches.append(cur)
        return batches

    best = cut(max_tokens)
    lo, hi = max(longest, -(-sum(lengths) // len(best))), max_tokens     # smallest cap that still needs len(best) batches
    while lo < hi:
        mid = (lo + hi) // 2
        b = cut(mid)
        if len(b) <= len(best):
            best, hi = b, mid
        else:
            lo = mid + 1
    return best


def fit_batch(workspace_bytes_fn, free_bytes: int, max_lookahead: int, fraction: float = 0.6, cap_tokens: int = 262144):
    """Largest (max_tokens, max_seqs) whose workspace fits into ``fraction`` of the free device memory: the KV cache of a
    multi-head 7B model is 0.5 MB per prompt token (32 layers x 2 x 4096 x 2 B), the cross-attention K/V of a T5-XL 0.2 MB, a
    multi-query Falcon's 8 KB -- one fixed token budget cannot suit all of them.  ``workspace_bytes_fn(tokens, seqs, lookahead)``
    is the library's own size query (b2s_workspace_bytes / b2s_encdec_workspace_bytes)."""
    budget = int(free_bytes * fraction)
    lo, hi = 2048, cap_tokens
    seqs = lambda t: max(64, t // 48)
    if workspace_bytes_fn(lo, seqs(lo), max_lookahead) > budget:
        raise _lib.B2SError(f"not enough free device memory for a 2048-token batch ({free_bytes >> 20} MiB free)")
    while lo < hi:
        mid = (lo + hi + 1) // 2
        if workspace_bytes_fn(mid, seqs(mid), max_lookahead) <= budget:
            lo = mid
        else:
            hi = mid - 1
    return lo, seqs(lo)


class Scorer:
    """One decoder resident on one GPU (the reference keeps one model resident at a time too,
    run_base_vs_instruct_100q.py:498-505).  ``max_tokens=None`` sizes the batch from the free device memory (:func:`fit_batch`)."""

    def __init__(self, shape: DecoderShape, weights_dev: Dict[str, torch.Tensor], mode: int = MODE_BF16_FAST,
                 device: Optional[torch.device] = None, max_tokens: Optional[int] = 65536, max_seqs: int = 1024,
                 max_lookahead: int = MAX_LOOK_AHEAD):
        if not torch.cuda.is_available():
            raise _lib.B2SError("no CUDA device: the B200 scorer has no CPU fallback")
        self.lib = _lib.load()
        self.shape = shape
        self.mode = mode
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.max_tokens, self.max_seqs, self.max_lookahead = max_tokens, int(max_seqs), int(max_lookahead)
        want = {MODE_FP32_CHECK: torch.float32, MODE_FP16: torch.float16, MODE_INT8_EMU: torch.float16}.get(mode, torch.bfloat16)
        W = dict(weights_dev)
        if mode == MODE_INT8_EMU:
            # bitsandbytes Linear8bitLt: every nn.Linear of the decoder layers is stored as int8 rows with one fp32 scale per output
            # row (lm_head is skipped, embeddings are not Linear); the library multiplies their dequantised form
            W = {k: (int8_rowwise_dequant(v) if k.split(".")[-1] in ("w_qkv", "w_dense", "w_up", "w_down") else v) for k, v in W.items()}
        if mode in (MODE_FP16, MODE_INT8_EMU):   # bf16 (or fp32) checkpoints are handed to the library as fp16 copies
            W = {k: (v.to(torch.float16) if _weights.is_matrix(k) and v.dtype != torch.float16 else v) for k, v in W.items()}
            if weights_dev.get("lm_head") is weights_dev.get("wte"):
                W["lm_head"] = W["wte"]
        mlp_type = shape.mlp_type
        if mlp_type == MLP_SWIGLU and mode != MODE_FP32_CHECK:
            # tensor-core modes: gate / up rows interleaved so that the projection's epilogue holds both halves of an output
            mlp_type = MLP_SWIGLU_FUSED
            ff = shape.d_ff
            for l in range(shape.n_layer):
                w = W[f"l{l}.w_up"]
                W[f"l{l}.w_up"] = torch.stack((w[:ff], w[ff:]), dim=1).reshape(2 * ff, w.shape[1]).contiguous()
        if "rope_cos" not in W and shape.rotary_dim > 0:
            W["rope_cos"], W["rope_sin"] = _weights.rope_tables(shape, self.device)
        for k, v in W.items():
            assert v.is_cuda and v.is_contiguous(), k
            if _weights.is_matrix(k):
                assert v.dtype == want, f"{k} is {v.dtype}, mode needs {want}"
            else:
                assert v.dtype == torch.float32, k
        self.W = W                      # keeps the tensors alive; the library borrows the pointers
        self._ptrs = _weights.pointer_table(shape, W)
        d = _lib.ModelDesc(abi_version=_lib.ABI_VERSION, arch=shape.arch, n_layer=shape.n_layer, d_model=shape.d_model,
                           n_head=shape.n_head, n_kv_head=shape.n_kv_head, head_dim=shape.head_dim, d_ff=shape.d_ff,
                           vocab=shape.vocab, rotary_dim=shape.rotary_dim, norm_eps=shape.norm_eps,
                           parallel_residual=int(shape.parallel_residual), n_ln_per_layer=shape.n_ln_per_layer,
                           has_bias=(1 if shape.has_bias else 2 if shape.qkv_bias else 0), qkv_layout=shape.qkv_layout, max_pos=shape.max_pos,
                           compute_mode=mode, norm_type=shape.norm_type, mlp_type=mlp_type,
                           pos_type=shape.pos_type, pos_offset=shape.pos_offset, act_type=shape.act_type,
                           emb_norm=int(shape.emb_norm))
        self._handle = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2s_model_create(C.byref(d), self._ptrs, len(self._ptrs), self.device.index or 0,
                                                 C.byref(self._handle)), "b2s_model_create")
        if self.max_tokens is None:      # batch sized from what is free on the device once the weights are resident
            torch.cuda.empty_cache()     # (cached blocks of a model that was just unloaded count as used otherwise)
            free, _ = torch.cuda.mem_get_info(self.device)
            self.max_tokens, self.max_seqs = fit_batch(lambda t, s, g: self.lib.b2s_workspace_bytes(self._handle, t, s, g), free,
                                                       self.max_lookahead)
        self.max_tokens = int(self.max_tokens)
        nbytes = self.lib.b2s_workspace_bytes(self._handle, self.max_tokens, self.max_seqs, self.max_lookahead)
        if nbytes == 0:
            raise _lib.B2SError("b2s_workspace_bytes returned 0")
        self.workspace = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        self.workspace_bytes = nbytes

    def close(self):
        if getattr(self, "_handle", None) is not None and self._handle:
            self.lib.b2s_model_destroy(self._handle)
            self._handle = None
        # drop the device memory this object keeps alive, so that it can be returned to the driver even while a caller
        # still holds the (closed) scorer
        self.workspace = None
        self.W = None
        self._ptrs = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- packing ----------------------------------------------------------------------------
    @staticmethod
    def pack(id_lists: Sequence[Sequence[int]], cand: np.ndarray, pin: bool = True):
        lens = np.fromiter((len(x) for x in id_lists), dtype=np.int64, count=len(id_lists))
        cu = np.zeros(len(id_lists) + 1, dtype=np.int32)
        np.cumsum(lens, out=cu[1:])
        toks = np.concatenate([np.asarray(x, dtype=np.int32) for x in id_lists]) if len(id_lists) else np.zeros(0, np.int32)
        t_tok, t_cu, t_cand = torch.from_numpy(toks), torch.from_numpy(cu), torch.from_numpy(np.ascontiguousarray(cand, dtype=np.int32))
        if pin:
            t_tok, t_cu, t_cand = t_tok.pin_memory(), t_cu.pin_memory(), t_cand.pin_memory()
        return t_tok, t_cu, t_cand, int(lens.max()) if len(lens) else 0

    # ---- scoring ----------------------------------------------------------------------------
    def score_batch_host(self, t_tok, t_cu, t_cand, topk: int, max_lookahead: int, eos_id: int,
                         out: torch.Tensor, gen: Optional[torch.Tensor], gen_tokens: Optional[int] = None):
        """One C-ABI call with host buffers (H2D + compute + D2H + sync inside).  With ``gen_tokens`` greedy decoding
        continues past the scan window (b2s_score_complete_host) and ``gen`` is [n, gen_tokens]."""
        n = t_cu.numel() - 1
        with torch.cuda.device(self.device):
            if gen_tokens is None or gen_tokens == max_lookahead:
                _lib.check(self.lib.b2s_score_host(self._handle, _lib.ptr(t_tok), _lib.ptr(t_cu), n, _lib.ptr(t_cand),
                                                   topk, max_lookahead, eos_id, _lib.ptr(out), _lib.ptr(gen),
                                                   _lib.ptr(self.workspace), self.workspace_bytes, _lib.current_stream()),
                           "b2s_score_host")
            else:
                _lib.check(self.lib.b2s_score_complete_host(self._handle, _lib.ptr(t_tok), _lib.ptr(t_cu), n, _lib.ptr(t_cand),
                                                            topk, max_lookahead, gen_tokens, eos_id, _lib.ptr(out), _lib.ptr(gen),
                                                            _lib.ptr(self.workspace), self.workspace_bytes,
                                                            _lib.current_stream()), "b2s_score_complete_host")

    def score_batch_device(self, d_tok, d_cu, d_cand, n: int, total: int, max_len: int, topk: int, max_lookahead: int,
                           eos_id: int, d_out: torch.Tensor, d_gen: Optional[torch.Tensor]):
        """One C-ABI call with device-resident inputs/outputs (enqueue only)."""
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2s_score(self._handle, _lib.ptr(d_tok), _lib.ptr(d_cu), n, total, max_len, _lib.ptr(d_cand),
                                          topk, max_lookahead, eos_id, _lib.ptr(d_out), _lib.ptr(d_gen),
                                          _lib.ptr(self.workspace), self.workspace_bytes, _lib.current_stream()),
                       "b2s_score")

    def score_ids(self, id_lists: Sequence[Sequence[int]], cand_pairs, topk: int = 5,
                  max_lookahead: int = MAX_LOOK_AHEAD, eos_id: Optional[int] = None, gen_tokens: Optional[int] = None,
                  nonfinite: str = "raise"):
        """Score already-tokenised prompts.  Returns (rows [N,8] float32, gen_ids [N,G] int32) in the callers' order.
        rows: p_a, p_b, rel, lse, logit_a, logit_b, pos_found, found.  G = ``max_lookahead``, or ``gen_tokens`` when
        given: the reference's max_new_tokens (50), generated for every prompt up to EOS while only the first
        ``max_lookahead`` positions are scanned (run_base_vs_instruct_100q.py:338-363).
        A prompt whose logits are not finite (an fp16 activation overflowed) never comes back as a number: with
        ``nonfinite="raise"`` a :class:`B2SRangeError` is raised once all batches ran, with ``"nan"`` its row is kept as
        NaN with found = -1 (the reference's NaN-row policy)."""
        N = len(id_lists)
        G = max_lookahead if gen_tokens is None else int(gen_tokens)
        if G < max_lookahead:
            raise ValueError(f"gen_tokens {G} < max_lookahead {max_lookahead}")
        if N == 0:   # empty input: nothing to do (the reference loop would simply not execute)
            return np.zeros((0, 8), np.float32), np.zeros((0, G), np.int32)
        eos = self.shape.eos_id if eos_id is None else int(eos_id)
        cand = np.asarray(cand_pairs, dtype=np.int32).reshape(-1, 2)
        if cand.shape[0] == 1 and N > 1:
            cand = np.repeat(cand, N, axis=0)
        assert cand.shape[0] == N
        V = self.shape.vocab
        if cand.min() < 0 or cand.max() >= V:
            raise ValueError("candidate id outside the vocabulary")
        lens = [len(x) for x in id_lists]
        if min(lens) <= 0:
            raise ValueError("empty prompt")
        if G > self.max_lookahead:
            raise ValueError(f"max_lookahead / gen_tokens {G} > workspace limit {self.max_lookahead}")
        rows = np.zeros((N, 8), np.float32)
        gens = np.full((N, G), -1, np.int32)
        range_error = None
        for batch in plan_batches(lens, self.max_tokens, self.max_seqs):
            t_tok, t_cu, t_cand, _ = self.pack([id_lists[i] for i in batch], cand[batch])
            out = torch.empty((len(batch), 8), dtype=torch.float32).pin_memory()
            gen = torch.empty((len(batch), G), dtype=torch.int32).pin_memory()
            try:
                self.score_batch_host(t_tok, t_cu, t_cand, topk, max_lookahead, eos, out, gen, gen_tokens=G)
            except _lib.B2SRangeError as e:      # rows were copied back; the affected ones are NaN with found = -1
                range_error = e
            rows[batch] = out.numpy()
            gens[batch] = gen.numpy()
        if range_error is not None and nonfinite == "raise":
            raise range_error
        return rows, gens


# ---- reference-shaped result dicts --------------------------------------------------------------
def _row_to_dict(row: np.ndarray, gen: np.ndarray, variant: str, prompt: str, tokenizer) -> dict:
    yes_prob, no_prob = float(row[0]), float(row[1])
    toks = [int(t) for t in gen if t >= 0]
    completion = tokenizer.decode(toks, skip_special_tokens=True) if tokenizer is not None else ""
    if variant == "100q":
        # run_base_vs_instruct_100q.py:370-382
        rel = yes_prob / (yes_prob + no_prob) if yes_prob + no_prob > 0 else 0.5
        return dict(yes_prob=yes_prob, no_prob=no_prob, relative_prob=rel, completion=completion.strip()[:100],
                    success=True)
    yes_lp = float(np.log(yes_prob)) if yes_prob > 0 else float("-inf")
    no_lp = float(np.log(no_prob)) if no_prob > 0 else float("-inf")
    out = dict(yes_logprob=yes_lp, no_logprob=no_lp, yes_prob=yes_prob, no_prob=no_prob)
    if variant == "instruct":   # compare_instruct_models.py:276-293
        out["relative_prob"] = yes_prob / (yes_prob + no_prob) if (yes_prob + no_prob) > 0 else float("nan")
    else:                       # compare_base_vs_instruct.py:290-305
        out["odds_ratio"] = yes_prob / no_prob if no_prob > 0 else float("inf")
    out.update(prompt=prompt, completion=completion.strip(), position_found=int(row[6]), yes_no_found=bool(row[7] > 0))
    return out


def _nan_dict(variant: str, prompt: str, err: Exception) -> dict:
    nan = float("nan")
    if variant == "100q":   # run_base_vs_instruct_100q.py:384-392
        return dict(yes_prob=nan, no_prob=nan, relative_prob=nan, completion=f"ERROR: {str(err)[:50]}", success=False)
    d = dict(yes_logprob=nan, no_logprob=nan, yes_prob=nan, no_prob=nan, prompt=prompt, completion=f"ERROR: {err}",
             position_found=-1, yes_no_found=False)
    d["relative_prob" if variant == "instruct" else "odds_ratio"] = nan
    return d


def candidate_ids_encdec(tokenizer, yes: str = "Yes", no: str = "No") -> Tuple[int, int]:
    """Candidate ids as the `is_encoder_decoder` branch of every reference script resolves them: first id of the bare
    word, special tokens included (compare_instruct_models.py:197-198, run_base_vs_instruct_100q.py:305-306,
    compare_base_vs_instruct.py:211-212; the T5 tokenizers append </s>, so [0] is the word piece)."""
    return int(tokenizer(yes).input_ids[0]), int(tokenizer(no).input_ids[0])


def tokenize_prompts(tokenizer, prompts, variant, cand_strings=None, encdec: bool = False):
    """Token ids per prompt and the [N,2] candidate-id table, resolved as the chosen entry point does (``encdec``: as its
    encoder-decoder branch does)."""
    ids = [list(tokenizer(p).input_ids) for p in prompts]
    resolve = (lambda *a: candidate_ids_encdec(tokenizer, *a)) if encdec else (lambda *a: candidate_ids(tokenizer, variant, *a))
    if cand_strings is None:
        cand = np.array([resolve()] * len(prompts), dtype=np.int32).reshape(-1, 2)
    else:
        cache = {}
        cand = np.zeros((len(prompts), 2), np.int32)
        for i, (a, b) in enumerate(cand_strings):
            if (a, b) not in cache:
                cache[(a, b)] = resolve(a, b)
            cand[i] = cache[(a, b)]
    return ids, cand


def score_prompts(scorer: Scorer, tokenizer, prompts: Sequence[str], variant: str = "100q",
                  cand_strings: Optional[Sequence[Tuple[str, str]]] = None,
                  max_lookahead: int = MAX_LOOK_AHEAD, shard_group=None, gen_tokens: Optional[int] = None) -> List[dict]:
    """Batched replacement of the reference's per-prompt loop.  ``gen_tokens`` (e.g. 50 = the reference's
    max_new_tokens) makes ``completion`` the full greedy continuation instead of the look-ahead window.  ``cand_strings`` optionally gives a
    (token_1, token_2) pair per prompt (the perturbation scenarios' ``target_tokens``); default is
    (" Yes", " No").  With ``shard_group`` (a torch.distributed group, or True for the default group) every rank
    scores its shard of the prompts and all ranks receive all rows (BASELINE config 4: sharded by prompt)."""
    V = VARIANTS[variant]
    try:
        ids, cand = tokenize_prompts(tokenizer, prompts, variant, cand_strings, encdec=getattr(scorer, "is_encoder_decoder", False))
        if shard_group is not None:
            from .dist import score_ids_sharded
            rows, gens = score_ids_sharded(scorer, ids, cand, topk=V["topk"], max_lookahead=max_lookahead,
                                           eos_id=tokenizer.eos_token_id, group=shard_group, gen_tokens=gen_tokens)
        else:
            rows, gens = scorer.score_ids(ids, cand, topk=V["topk"], max_lookahead=max_lookahead,
                                          eos_id=tokenizer.eos_token_id, gen_tokens=gen_tokens, nonfinite="nan")
    except Exception as e:   # reference error policy: NaN rows, sweep continues
        if variant != "100q":
            raise
        return [_nan_dict(variant, p, e) for p in prompts]
    # found = -1: the logits of this prompt were not finite (fp16 overflow) -- the reference's NaN row, never a number
    return [_row_to_dict(rows[i], gens[i], variant, prompts[i], tokenizer) if rows[i][7] >= 0
            else _nan_dict(variant, prompts[i], ArithmeticError("non-finite logits (activation overflow in the 16-bit mode)"))
            for i in range(len(prompts))]


def get_yes_no_logprobs(model: Scorer, tokenizer, prompt: str, device=None, variant: str = "100q") -> dict:
    """Drop-in for the reference function of the same name (one prompt)."""
    return score_prompts(model, tokenizer, [prompt], variant=variant)[0]
