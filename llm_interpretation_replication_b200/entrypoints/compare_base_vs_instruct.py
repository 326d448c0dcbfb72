#!/usr/bin/env python
"""Entry point 6 -- mirrors reference analysis/compare_base_vs_instruct.py, the script that produced
``data/model_comparison_results.csv`` (which entry points 3, 4 and 5 consume).

Kept from the reference: the nine active (base, instruct) pairs (:136-180); the prompt formats chosen BY MODEL NAME
(:462-463) -- few-shot + ``Question: ... \\nAnswer:`` when "base" occurs in the model name or the model is
``bigscience/bloom-7b1``, otherwise few-shot + the bare question (so ``tiiuae/falcon-7b``, ``EleutherAI/pythia-6.9b``,
``meta-llama/Llama-2-7b-hf`` ... get the second form although they are base models: preserved); the "base_vs" scorer
variant (top-2, candidate ids WITH special tokens, ``max_length = len + 50``, ``odds_ratio = yes/no`` or inf; :185-305);
NaN rows for a model that cannot be run or a prompt that fails (:472-492); ``prepare_data_for_csv`` (:90-111) with the
columns prompt, model, model_family, base_or_instruct, model_output, yes_prob, no_prob, odds_ratio and
``model_family = name.split('/')[1].split('-')[0].lower()``.  The per-prompt loop (:456-480) is one batched call into
libb200score per model.  The T5 pair runs through ``t5.EncDecScorer`` (the script's encoder-decoder branch :192-237: bare-word
candidate ids, ``max_length=50`` = 49 generated tokens), the Qwen pair through the Llama path with a q|k|v bias; a model
outside the supported families produces the reference's NaN rows.  The reference's cache purge (:62-88) is deliberately not reproduced.

    python -m llm_interpretation_replication_b200.entrypoints.compare_base_vs_instruct [--out DIR] [--layers N] [--questions K]
"""
from __future__ import annotations

import argparse
import os

import pandas as pd
import torch

from .. import scorer as S
from .common import COMPLETION_TOKENS, FEW_SHOT, load_model, log_print, maybe_init_dist, survey_questions, unload

# reference :136-180 (the commented-out pairs -- MPT, Baichuan2, XGen -- are inactive there too)
model_pairs = [
    ("google/t5-v1_1-base", "google/flan-t5-base"),
    ("EleutherAI/pythia-6.9b", "databricks/dolly-v2-7b"),
    ("stabilityai/stablelm-base-alpha-7b", "stabilityai/stablelm-tuned-alpha-7b"),
    ("meta-llama/Llama-2-7b-hf", "meta-llama/Llama-2-7b-chat-hf"),
    ("tiiuae/falcon-7b", "tiiuae/falcon-7b-instruct"),
    ("mistralai/Mistral-7B-v0.1", "mistralai/Mistral-7B-Instruct-v0.2"),
    ("Qwen/Qwen-7B", "Qwen/Qwen-7B-Chat"),
    ("togethercomputer/RedPajama-INCITE-7B-Base", "togethercomputer/RedPajama-INCITE-7B-Instruct"),
    ("bigscience/bloom-7b1", "bigscience/bloomz-7b1"),
]


def format_prompt(model_name: str, prompt: str) -> str:
    """reference :462-463, including its name test."""
    if "base" in model_name.lower() or model_name.lower() in ["bigscience/bloom-7b1"]:
        return f"{FEW_SHOT}Question: {prompt} Answer either 'Yes' or 'No', without any other text.\nAnswer:"
    return f"{FEW_SHOT}{prompt} Answer either 'Yes' or 'No', without any other text."


def prepare_data_for_csv(outputs, pairs):
    """reference :90-111."""
    data = []
    for base_name, instruct_name in pairs:
        for model_name in [base_name, instruct_name]:
            model_family = model_name.split("/")[1].split("-")[0].lower()
            base_or_instruct = "base" if model_name == base_name else "instruct"
            for prompt in outputs.get(model_name, {}):
                result = outputs[model_name][prompt]
                data.append({"prompt": prompt, "model": model_name, "model_family": model_family,
                             "base_or_instruct": base_or_instruct, "model_output": result.get("completion", "N/A"),
                             "yes_prob": result.get("yes_prob", float("nan")), "no_prob": result.get("no_prob", float("nan")),
                             "odds_ratio": result.get("odds_ratio", float("nan"))})
    return pd.DataFrame(data)


def run(prompts, pairs=None, n_layer=0, device=None, shard_group=None):
    outputs = {}
    for base_name, instruct_name in (pairs or model_pairs):
        for model_name in [base_name, instruct_name]:
            outputs[model_name] = {}
            scorer = None
            formatted = [format_prompt(model_name, p) for p in prompts]
            try:
                scorer, tok = load_model(model_name, device=device, n_layer=n_layer)
                # reference :199 generates with max_length=50 in the encoder-decoder branch: the start token counts, 49 new ones
                n_gen = COMPLETION_TOKENS - 1 if getattr(scorer, "is_encoder_decoder", False) else COMPLETION_TOKENS
                rows = S.score_prompts(scorer, tok, formatted, variant="base_vs", shard_group=shard_group, gen_tokens=n_gen)
                for p, r in zip(prompts, rows):
                    outputs[model_name][p] = r
            except Exception as e:      # reference :482-492: the error is stored for every prompt of the model
                log_print(f"Error running {model_name}: {e}")
                for p, f in zip(prompts, formatted):
                    outputs[model_name][p] = {"yes_prob": float("nan"), "no_prob": float("nan"), "odds_ratio": float("nan"),
                                              "prompt": f, "completion": "N/A"}
            finally:
                if scorer is not None:
                    unload(scorer)
    return outputs


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.environ.get("B2S_OUTPUT_DIR", "llm_interpretation_out"))
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--questions", type=int, default=0)
    ap.add_argument("--pairs", nargs="*", default=None, help="subset of the pair list, by base-model name")
    args = ap.parse_args(argv)
    rank, world, group = maybe_init_dist()
    os.makedirs(args.out, exist_ok=True)
    prompts, _ = survey_questions()          # the script's 50 questions are the first survey's list (:308-359)
    if args.questions:
        prompts = prompts[:args.questions]
    pairs = [p for p in model_pairs if args.pairs is None or p[0] in args.pairs]
    dev = f"cuda:{torch.cuda.current_device()}" if torch.cuda.is_available() else None
    outputs = run(prompts, pairs=pairs, n_layer=args.layers, device=dev, shard_group=group)
    df = prepare_data_for_csv(outputs, pairs)
    if rank != 0:
        return df
    path = os.path.join(args.out, "model_comparison_results.csv")
    df.to_csv(path, index=False)
    log_print(f"Results saved to {path}")
    return df


if __name__ == "__main__":
    main()
