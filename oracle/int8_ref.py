"""bitsandbytes LLM.int8 (``Linear8bitLt`` / ``MatMul8bitLt``), restated on the CPU -- test infrastructure only.

The reference's GPU configuration loads every local model with ``BitsAndBytesConfig(load_in_8bit=True)``
(analysis/run_base_vs_instruct_100q.py:84,433-437 ``USE_8BIT``; compare_instruct_models.py:412-416).  bitsandbytes is a
third-party dependency that is absent from /root/reference and from this image (requirements: ``bitsandbytes``, unpinned), so this
file restates its published algorithm (Dettmers et al., "LLM.int8()", 2022; bitsandbytes/autograd/_functions.py MatMul8bitLt.forward,
bitsandbytes/functional.py double_quant / igemmlt / mm_dequant) and **parity against the library itself is unpinned**:

  load time   CB[n, :] = rint(W[n, :] * 127 / SCB[n]),  SCB[n] = max |W[n, :]|                       (int8 rows + fp32 row scales)
  forward, for the rows A of ONE call (one prompt's tokens, or one token in a cached step), fp16:
    outlier columns   idx = {k : any |A[m, k]| >= threshold}                                          (threshold 6.0)
    row scales        SCA[m] = max {|A[m, k]| : |A[m, k]| < threshold}       (outlier ELEMENTS do not enter the statistics)
    int8 rows         CA[m, k] = rint(A[m, k] * 127 / SCA[m]),  CA[:, idx] = 0
    int32 product     out32 = CA . CB^T ;  out = out32 * SCA[m] * SCB[n] / 127^2                      (+ bias), cast to fp16
    fp16 part         out += A[:, idx] . (CB[:, idx] * SCB / 127)^T                                  (dequantised weight columns)
  lm_head is skipped (``llm_int8_skip_modules`` default), embeddings are not Linear.

The integer product is formed exactly here (float64 holds every partial sum of |.| <= 127^2 * K).  The CUDA path multiplies the
same grid values as fp16 x fp16 -> fp32 (B2S_MODE_INT8_EMU); what separates the two is the fp16 rounding of the dequantised
operands and the fp32 accumulation order.
"""
from __future__ import annotations

import torch

THRESHOLD = 6.0


def _div127(s: torch.Tensor) -> torch.Tensor:
    """fp32(127 / s), correctly rounded (``127.0 / tensor`` is reciprocal-then-multiply in torch: one ulp off now and then,
    which moves exact rounding ties)."""
    return torch.full_like(s, 127.0) / s


def quantize_weight(w: torch.Tensor):
    """(CB int8-valued float tensor, SCB row scales) of a weight matrix [N, K]."""
    wf = w.float()
    scb = wf.abs().amax(dim=1).clamp_min(1e-30)
    cb = torch.round(wf * _div127(scb)[:, None])
    return cb, scb


def linear8bit(a: torch.Tensor, w: torch.Tensor, bias=None, threshold: float = THRESHOLD, out_fp16: bool = True) -> torch.Tensor:
    """One ``Linear8bitLt`` call on the rows ``a`` [M, K] (already fp16-representable) with weight ``w`` [N, K]."""
    cb, scb = quantize_weight(w)
    af = a.float()
    big = af.abs() >= threshold if threshold > 0 else torch.zeros_like(af, dtype=torch.bool)
    idx = big.any(dim=0)                                         # outlier columns of this call
    sca = torch.where(big, torch.zeros_like(af), af.abs()).amax(dim=1)
    safe = sca.clamp_min(1e-30)
    ca = torch.round(af * _div127(safe)[:, None])
    ca = torch.where(big, torch.zeros_like(ca), ca)
    ca[:, idx] = 0
    out32 = ca.double() @ cb.double().t()                        # exact integers
    out = (out32 * (sca.double()[:, None] * scb.double()[None, :]) / (127.0 * 127.0)).float()
    if bias is not None:
        out = out + bias.float()
    if out_fp16:
        out = out.to(torch.float16).float()
    if idx.any():
        sub_b = (cb[:, idx] * scb[:, None] / 127.0).to(torch.float16).float()
        out = out + af[:, idx] @ sub_b.t()
        if out_fp16:
            out = out.to(torch.float16).float()
    return out


def make_linear_fn(round_act, prompt_len: int, threshold: float = THRESHOLD, out_fp16: bool = False):
    """``linear_fn`` for oracle/decoder_ref.forward_hidden: every block's linear layer becomes Linear8bitLt calls on the
    fp16-rounded input rows.  ``generate`` with a KV cache calls each layer once on the prompt's rows and then once per new
    token; the cache-free restatement recomputes all rows at every position, so the rows are cut the same way here -- rows
    [0, prompt_len) are one call, every later row its own -- which reproduces the cached values (outlier columns are a
    property of a call's rows)."""
    def fn(x, w, b):
        parts = [linear8bit(round_act(x[:prompt_len]), w, b, threshold=threshold, out_fp16=out_fp16)]
        for r in range(prompt_len, x.shape[0]):
            parts.append(linear8bit(round_act(x[r:r + 1]), w, b, threshold=threshold, out_fp16=out_fp16))
        return torch.cat(parts, dim=0)
    return fn
