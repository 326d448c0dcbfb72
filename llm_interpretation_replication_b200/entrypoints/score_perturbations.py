#!/usr/bin/env python
"""Entry point 7 -- BASELINE config 3 on the real strings: every rephrasing of ``data/perturbations.json`` scored by a
local decoder.

The reference ran this experiment through hosted APIs (analysis/perturb_prompts.py); what is kept here is everything
on that path that is not the HTTP client:
  * the prompt template ``f"{rephrased_main} {orig_format}"`` (perturb_prompts.py:215, "binary" format);
  * one candidate pair per scenario = its ``target_tokens`` (:201, :485-488).  The API matched the bare strings against
    the returned top-logprob tokens; a local decoder needs token ids, resolved the way the reference's local scorer
    resolves " Yes"/" No": first id of ``tokenizer(" " + token, add_special_tokens=False)``
    (run_base_vs_instruct_100q.py:331-335) -- a stated choice, the reference never ran this set locally;
  * the result row schema (:529-545): Model, Original Main Part, Response Format, Confidence Format, Rephrased Main
    Part, Full Rephrased Prompt, Model Response, Token_1_Prob, Token_2_Prob, Odds_Ratio (``inf`` when Token_2_Prob is
    0, :521), plus ``Relative_Prob`` exactly as the analysis derives it (analyze_perturbation_results.py:1738-1746:
    Token_1_Prob / (Token_1_Prob + Token_2_Prob), NaN when the sum is 0).
Scoring follows the local-decoder semantics of run_base_vs_instruct_100q.py (top-5, 10-position look-ahead, position-0
fallback).  The 10,000 prompts are one length-bucketed batched run (sharded by prompt under torchrun); duplicate
rephrasings (60/23/28/19/48 per scenario) are scored like any other prompt and come out identical.

    python -m llm_interpretation_replication_b200.entrypoints.score_perturbations [--model NAME] [--data PATH] [--out DIR]
"""
from __future__ import annotations

import argparse
import gzip
import json
import os

import numpy as np
import pandas as pd
import torch

from .. import scorer as S
from .common import COMPLETION_TOKENS, DATA_DIR, format_perturbation, load_model, log_print, maybe_init_dist, unload

RESULTS_NAME = "perturbation_results_local.csv"


def load_perturbations(path: str | None = None):
    """[{original_main, response_format, target_tokens, confidence_format, rephrasings}] x 5 (the file the reference's
    perturb_prompts.py writes and reads, :966-969).  Default: the copy shipped next to the package (gzip)."""
    path = path or os.environ.get("B2S_PERTURBATIONS") or os.path.join(DATA_DIR, "perturbations.json.gz")
    opener = gzip.open if path.endswith(".gz") else open
    with opener(path, "rt", encoding="utf-8") as f:
        return json.load(f)


def build_requests(scenarios, limit: int = 0):
    """(prompt, (cand_1, cand_2), scenario index, rephrasing) in file order; ``limit`` keeps the first k per scenario."""
    reqs = []
    for si, sc in enumerate(scenarios):
        t1, t2 = sc["target_tokens"]
        reph = sc["rephrasings"][:limit] if limit else sc["rephrasings"]
        for r in reph:
            reqs.append((format_perturbation(r, sc["response_format"]), (" " + t1, " " + t2), si, r))
    return reqs


def run(scenarios, model_name="tiiuae/falcon-7b", n_layer=0, device=None, shard_group=None, limit=0, max_tokens=262144):
    reqs = build_requests(scenarios, limit)
    scorer, tok = load_model(model_name, device=device, n_layer=n_layer, max_tokens=max_tokens, max_seqs=4096)
    try:
        for sc in scenarios:      # the offline hash tokenizer must not map a pair to one id
            a, b = S.candidate_ids(tok, "100q", " " + sc["target_tokens"][0], " " + sc["target_tokens"][1])
            if a == b:
                raise ValueError(f"target tokens {sc['target_tokens']} resolve to the same id {a}")
        rows = S.score_prompts(scorer, tok, [q[0] for q in reqs], variant="100q", cand_strings=[q[1] for q in reqs],
                               shard_group=shard_group, gen_tokens=COMPLETION_TOKENS)
    finally:
        unload(scorer)
    out = []
    for (prompt, _, si, reph), r in zip(reqs, rows):
        sc = scenarios[si]
        t1p, t2p = r["yes_prob"], r["no_prob"]
        out.append({"Model": model_name, "Original Main Part": sc["original_main"], "Response Format": sc["response_format"],
                    "Confidence Format": sc["confidence_format"], "Rephrased Main Part": reph, "Full Rephrased Prompt": prompt,
                    "Model Response": r["completion"], "Token_1_Prob": t1p, "Token_2_Prob": t2p,
                    "Odds_Ratio": (t1p / t2p if t2p > 0 else float("inf")) if np.isfinite(t1p) else float("nan")})
    df = pd.DataFrame(out)
    total = df["Token_1_Prob"] + df["Token_2_Prob"]          # analyze_perturbation_results.py:1738-1746
    df["Total_Prob"] = total
    df["Relative_Prob"] = np.where(total > 0, df["Token_1_Prob"] / total, np.nan)
    return df


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="tiiuae/falcon-7b")
    ap.add_argument("--data", default=None, help="perturbations.json (or .json.gz); default: the shipped copy")
    ap.add_argument("--out", default=os.environ.get("B2S_OUTPUT_DIR", "llm_interpretation_out"))
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--limit", type=int, default=0, help="first k rephrasings of every scenario only")
    args = ap.parse_args(argv)
    rank, world, group = maybe_init_dist()
    os.makedirs(args.out, exist_ok=True)
    dev = f"cuda:{torch.cuda.current_device()}" if torch.cuda.is_available() else None
    df = run(load_perturbations(args.data), model_name=args.model, n_layer=args.layers, device=dev, shard_group=group,
             limit=args.limit)
    if rank != 0:
        return df
    path = os.path.join(args.out, RESULTS_NAME)
    df.to_csv(path, index=False)
    log_print(f"Results saved to {path} ({len(df)} rows)")
    return df


if __name__ == "__main__":
    main()
