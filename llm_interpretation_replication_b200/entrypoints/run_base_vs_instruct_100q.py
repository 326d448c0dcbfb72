#!/usr/bin/env python
"""Entry point 1 -- mirrors reference analysis/run_base_vs_instruct_100q.py.

Same constants, prompt formats, checkpoint JSON (`{'completed_models': [...], 'results': [row dicts]}`, :265-276),
result CSV columns (yes_prob, no_prob, relative_prob, completion, success, prompt, model, formatted_prompt,
model_family, base_or_instruct; :376-382,473-475,547-549) and `analyze_results(df)` (:606-712).  The per-prompt
`get_yes_no_logprobs` loop (:464-476) is one batched call into libb200score; the bootstrap loops (:655-658,
:692-695) are the device kernels.  Models are loaded from a local checkpoint directory when one is given (checkpoint.py), else random-init stand-ins of the
named hub shapes (no hub access here).

    python -m llm_interpretation_replication_b200.entrypoints.run_base_vs_instruct_100q [--out DIR] [--layers N] [--questions K]
"""
from __future__ import annotations

import argparse
import json
import os

import numpy as np
import pandas as pd

from .. import bootstrap, scorer as S
from .common import COMPLETION_TOKENS, format_base, format_instruct, load_model, log_print, survey_questions, unload

OUTPUT_DIR = os.environ.get("B2S_OUTPUT_DIR", "llm_interpretation_out")
CHECKPOINT_NAME = "base_vs_instruct_100q_checkpoint.json"
RESULTS_NAME = "base_vs_instruct_100q_results.csv"

# reference :88-103: all six active pairs (LLaMA-2 and Gemma are commented out there too)
MODEL_PAIRS = [
    ("tiiuae/falcon-7b", "tiiuae/falcon-7b-instruct", "Falcon"),
    ("stabilityai/stablelm-base-alpha-7b", "stabilityai/stablelm-tuned-alpha-7b", "StableLM"),
    ("togethercomputer/RedPajama-INCITE-7B-Base", "togethercomputer/RedPajama-INCITE-7B-Instruct", "RedPajama"),
    ("bigscience/bloom-7b1", "bigscience/bloomz-7b1", "BLOOM"),
    ("EleutherAI/pythia-6.9b", "databricks/dolly-v2-7b", "Pythia-Dolly"),
    ("mistralai/Mistral-7B-v0.1", "mistralai/Mistral-7B-Instruct-v0.2", "Mistral"),
]


def load_checkpoint(path):
    if os.path.exists(path):
        with open(path, "r") as f:
            return json.load(f)
    return {"completed_models": [], "results": []}


def save_checkpoint(path, checkpoint):
    with open(path, "w") as f:
        json.dump(checkpoint, f, indent=2)


def run_model_on_prompts(model_name, prompts, is_base_model=True, n_layer=0, device=None):
    """reference :395-507 -- same result rows, same NaN rows when the model cannot be run."""
    log_print(f"\n{'=' * 60}\nLoading model: {model_name}")
    results = []
    scorer = None
    try:
        scorer, tok = load_model(model_name, device=device, n_layer=n_layer)
        formatted = [format_base(p) if is_base_model else format_instruct(p) for p in prompts]
        rows = S.score_prompts(scorer, tok, formatted, variant="100q", gen_tokens=COMPLETION_TOKENS)
        for p, f, r in zip(prompts, formatted, rows):
            r["prompt"] = p
            r["model"] = model_name
            r["formatted_prompt"] = f[:200]
            results.append(r)
        log_print(f"Finished {model_name}. Processed {len(results)} prompts.")
    except Exception as e:
        log_print(f"ERROR loading/running {model_name}: {e}")
        results = [{"prompt": p, "model": model_name, "yes_prob": float("nan"), "no_prob": float("nan"),
                    "relative_prob": float("nan"), "completion": f"MODEL_ERROR: {str(e)[:50]}", "success": False}
                   for p in prompts]
    finally:
        if scorer is not None:
            unload(scorer)
    return results


def analyze_results(df):
    """reference :606-712 -- per-family and overall bootstrap of mean(instruct - base), one seed for all."""
    fam_diffs = {}
    for family in df["model_family"].unique():
        fdf = df[df["model_family"] == family]
        merged = pd.merge(fdf[fdf["base_or_instruct"] == "base"][["prompt", "relative_prob"]],
                          fdf[fdf["base_or_instruct"] == "instruct"][["prompt", "relative_prob"]],
                          on="prompt", suffixes=("_base", "_instruct")).dropna()
        if len(merged) < 10:
            log_print(f"{family}: Insufficient data ({len(merged)} valid pairs)")
            continue
        fam_diffs[family] = merged["relative_prob_instruct"].values - merged["relative_prob_base"].values
    results, overall = bootstrap.analyze_results_bootstrap(fam_diffs)
    for family, r in results.items():
        p = r["p_value"]
        stars = "***" if p < 0.01 else ("**" if p < 0.05 else ("*" if p < 0.1 else ""))
        log_print(f"{family}: Mean Diff = {r['mean_diff']:+.3f} [{r['ci_lower']:+.3f}, {r['ci_upper']:+.3f}] {stars} (n={r['n_prompts']})")
    if overall is not None:
        log_print(f"\nOVERALL (across {len(results)} model families):")
        log_print(f"  Mean Diff = {overall['mean_diff']:+.3f} [{overall['ci_lower']:+.3f}, {overall['ci_upper']:+.3f}]")
        log_print(f"  MAE = {overall['mae']:.3f}\n  P-value = {overall['p_value']:.4f}")
    return results


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=OUTPUT_DIR)
    ap.add_argument("--layers", type=int, default=0, help="truncate every model to N layers (smoke runs)")
    ap.add_argument("--questions", type=int, default=0, help="use only the first K questions (smoke runs)")
    ap.add_argument("--pairs", type=int, default=0, help="use only the first P model pairs")
    args = ap.parse_args(argv)
    os.makedirs(args.out, exist_ok=True)
    ckpt_path = os.path.join(args.out, CHECKPOINT_NAME)
    s1, s2 = survey_questions()
    prompts = s1 + s2
    if args.questions:
        prompts = prompts[:args.questions]
    pairs = MODEL_PAIRS[:args.pairs] if args.pairs else MODEL_PAIRS
    checkpoint = load_checkpoint(ckpt_path)
    completed = set(checkpoint["completed_models"])
    all_results = checkpoint["results"]
    for base_model, instruct_model, family in pairs:
        for name, is_base in ((base_model, True), (instruct_model, False)):
            if name in completed:
                log_print(f"Skipping {name} (already completed)")
                continue
            rows = run_model_on_prompts(name, prompts, is_base_model=is_base, n_layer=args.layers)
            for r in rows:
                r["model_family"] = family
                r["base_or_instruct"] = "base" if is_base else "instruct"
            all_results.extend(rows)
            checkpoint["completed_models"].append(name)
            checkpoint["results"] = all_results
            save_checkpoint(ckpt_path, checkpoint)
            log_print(f"Checkpoint saved after {name}")
    df = pd.DataFrame(all_results)
    out_csv = os.path.join(args.out, RESULTS_NAME)
    df.to_csv(out_csv, index=False)
    log_print(f"Results saved to: {out_csv}\nTotal rows: {len(df)}")
    analyze_results(df)
    return df


if __name__ == "__main__":
    main()
