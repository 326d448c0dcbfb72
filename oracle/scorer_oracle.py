"""Restatement of the reference's ``get_yes_no_logprobs`` decoder-only branch, all three script
variants (test infrastructure only; see oracle/__init__.py).

Variant table = SURVEY.md Appendix A:
  "100q"      analysis/run_base_vs_instruct_100q.py:279-392     top-k 5, candidates without specials,
                                                                 rel fallback 0.5, completion[:100]
  "instruct"  analysis/compare_instruct_models.py:171-293        top-k 2, candidates WITH specials,
                                                                 rel fallback NaN, logprobs, position_found
  "base_vs"   analysis/compare_base_vs_instruct.py:185-305       as "instruct" but max_length=len+50 and
                                                                 odds_ratio = yes/no (inf if no == 0)
The transformer arithmetic is HF ``model.generate`` itself (the reference's own third-party
dependency, ``transformers`` 5.5.0 here).  Pinned by tests/golden/scorer_golden.json, produced by
calling the reference's function object imported from /root/reference.
"""
from __future__ import annotations

import math
from typing import Optional

import numpy as np
import torch

MAX_LOOK_AHEAD = 10

VARIANTS = {
    "100q": dict(topk=5, cand_specials=False, rel_fallback=0.5),
    "instruct": dict(topk=2, cand_specials=True, rel_fallback=float("nan")),
    "base_vs": dict(topk=2, cand_specials=True, rel_fallback=None),
}


def candidate_ids(tokenizer, variant: str, yes=" Yes", no=" No"):
    """Candidate-token resolution exactly as each entry point does it."""
    if VARIANTS[variant]["cand_specials"]:
        y = tokenizer(yes).input_ids
        n = tokenizer(no).input_ids
    else:
        y = tokenizer(yes, add_special_tokens=False).input_ids
        n = tokenizer(no, add_special_tokens=False).input_ids
    y_id = y[0] if y else tokenizer.encode(yes.strip())[0]
    n_id = n[0] if n else tokenizer.encode(no.strip())[0]
    return int(y_id), int(n_id)


def get_yes_no_logprobs(model, tokenizer, prompt: str, device="cpu", variant: str = "100q",
                        cand_ids: Optional[tuple] = None, max_new_tokens: int = 50):
    V = VARIANTS[variant]
    inputs = tokenizer(prompt, return_tensors="pt").to(device)
    yes_id, no_id = cand_ids if cand_ids is not None else candidate_ids(tokenizer, variant)
    if variant == "100q":
        pad = tokenizer.pad_token_id if tokenizer.pad_token_id else tokenizer.eos_token_id
    elif variant == "instruct":
        pad = tokenizer.pad_token_id if tokenizer.pad_token_id is not None else tokenizer.eos_token_id
    else:
        pad = tokenizer.pad_token_id
    kw = dict(num_return_sequences=1, output_scores=True, return_dict_in_generate=True,
              pad_token_id=pad, eos_token_id=tokenizer.eos_token_id)
    if variant == "base_vs":
        kw["max_length"] = len(inputs.input_ids[0]) + max_new_tokens
    else:
        kw["max_new_tokens"] = max_new_tokens
    with torch.no_grad():
        outputs = model.generate(inputs.input_ids, **kw)

    found, position = False, 0
    for pos, scores in enumerate(outputs.scores[:MAX_LOOK_AHEAD]):
        probs = torch.softmax(scores[0], dim=-1)
        _, top_tokens = torch.topk(probs, k=V["topk"])
        if yes_id in top_tokens or no_id in top_tokens:
            yes_prob = probs[yes_id].item()
            no_prob = probs[no_id].item()
            found, position = True, pos
            break
    if not found:
        probs = torch.softmax(outputs.scores[0][0], dim=-1)
        yes_prob = probs[yes_id].item()
        no_prob = probs[no_id].item()
        position = 0
    gen = outputs.sequences[0][len(inputs.input_ids[0]):]
    completion = tokenizer.decode(gen, skip_special_tokens=True)

    if variant == "100q":
        rel = yes_prob / (yes_prob + no_prob) if yes_prob + no_prob > 0 else 0.5
        return dict(yes_prob=yes_prob, no_prob=no_prob, relative_prob=rel,
                    completion=completion.strip()[:100], success=True,
                    _position_found=position, _found=found, _gen_ids=[int(t) for t in gen],
                    _n_scores=len(outputs.scores))
    yes_lp = float(np.log(yes_prob)) if yes_prob > 0 else float("-inf")
    no_lp = float(np.log(no_prob)) if no_prob > 0 else float("-inf")
    out = dict(yes_logprob=yes_lp, no_logprob=no_lp, yes_prob=yes_prob, no_prob=no_prob)
    if variant == "instruct":
        out["relative_prob"] = yes_prob / (yes_prob + no_prob) if (yes_prob + no_prob) > 0 else np.nan
    else:
        out["odds_ratio"] = yes_prob / no_prob if no_prob > 0 else float("inf")
    out.update(prompt=prompt, completion=completion.strip(), position_found=position, yes_no_found=found,
               _gen_ids=[int(t) for t in gen], _n_scores=len(outputs.scores))
    return out
