#!/usr/bin/env python
"""Entry point 2 -- mirrors reference analysis/compare_instruct_models.py and ..._survey2.py (byte-identical in
the reference except for the prompt list and the CSV name).

Keeps the model list (:145-166), the bare / Baichuan prompt formats (:488-492), the k=2 scorer variant with
candidate ids resolved WITH special tokens (:229-233), the NaN rows for models that cannot be run (:502-522), and
`prepare_data_for_csv` (:103-121) with columns prompt, model, model_family, model_output, yes_prob, no_prob,
relative_prob.  All ten models run: the Falcon, GPT-NeoX, Mistral, OPT and BLOOM families through the decoder-only
scorer, the encoder-decoder models (tk-instruct, T0: the T5 branch at :178-226) through ``t5.EncDecScorer`` with the bare-word
candidate ids of that branch (:197-198), and the two whose hub repositories ship custom modelling code (Baichuan2-7B: Llama block
with packed W_pack and a row-normalised head; Qwen-7B: Llama block with a q|k|v bias) through the Llama path -- their math is
pinned against HF's Qwen2 / Llama classes, their checkpoint key maps follow the published code and are not pinnable offline
(``checkpoint.py``).  A model outside these families gets the reference's NaN rows.  The reference's cache purge (:81-101) is deliberately not reproduced.

    python -m llm_interpretation_replication_b200.entrypoints.compare_instruct_models [--survey 1|2] [--out DIR]
"""
from __future__ import annotations

import argparse
import os

import pandas as pd

from .. import scorer as S
import torch

from .common import COMPLETION_TOKENS, format_baichuan, format_instruct, load_model, log_print, maybe_init_dist, survey_questions, unload

models = [
    "allenai/tk-instruct-3b-def",
    "baichuan-inc/Baichuan2-7B-Chat",
    "bigscience/bloomz-7b1",
    "bigscience/T0_3B",
    "facebook/opt-iml-1.3b",
    "h2oai/h2ogpt-oasst1-512-12b",
    "mistralai/Mistral-7B-Instruct-v0.3",
    "Qwen/Qwen-7B-Chat",
    "tiiuae/falcon-7b-instruct",
    "togethercomputer/RedPajama-INCITE-7B-Instruct",
]


def prepare_data_for_csv(outputs, model_names):
    data = []
    for model_name in model_names:
        model_family = model_name.split("/")[1].split("-")[0].lower()
        for prompt in outputs.get(model_name, {}):
            result = outputs[model_name][prompt]
            data.append({"prompt": prompt, "model": model_name, "model_family": model_family,
                         "model_output": result.get("completion", "N/A"),
                         "yes_prob": result.get("yes_prob", float("nan")),
                         "no_prob": result.get("no_prob", float("nan")),
                         "relative_prob": result.get("relative_prob", float("nan"))})
    return pd.DataFrame(data)


def run(prompts, model_names=None, n_layer=0, device=None, shard_group=None):
    """With shard_group (torchrun, one rank per GPU) the prompts of every model are sharded across the ranks and
    the rows all-gathered (BASELINE config 4); every rank returns the full table."""
    outputs = {}
    for model_name in (model_names or models):
        outputs[model_name] = {}
        scorer = None
        try:
            scorer, tok = load_model(model_name, device=device, n_layer=n_layer)
            fmt = format_baichuan if "baichuan" in model_name.lower() else format_instruct
            formatted = [fmt(p) for p in prompts]
            rows = S.score_prompts(scorer, tok, formatted, variant="instruct", shard_group=shard_group,
                                   gen_tokens=COMPLETION_TOKENS)
            for p, r in zip(prompts, rows):
                outputs[model_name][p] = r
        except Exception as e:
            log_print(f"Error running {model_name}: {e}")
            for p in prompts:
                outputs[model_name][p] = {"yes_prob": float("nan"), "no_prob": float("nan"), "relative_prob": float("nan"),
                                          "prompt": p, "completion": "N/A"}
        finally:
            if scorer is not None:
                unload(scorer)
    return outputs


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--survey", type=int, default=1, choices=[1, 2])
    ap.add_argument("--out", default=os.environ.get("B2S_OUTPUT_DIR", "llm_interpretation_out"))
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--questions", type=int, default=0)
    ap.add_argument("--models", nargs="*", default=None, help="subset of the model list")
    args = ap.parse_args(argv)
    rank, world, group = maybe_init_dist()
    os.makedirs(args.out, exist_ok=True)
    s1, s2 = survey_questions()
    prompts = s1 if args.survey == 1 else s2
    if args.questions:
        prompts = prompts[:args.questions]
    names = args.models or models
    dev = f"cuda:{torch.cuda.current_device()}" if torch.cuda.is_available() else None
    outputs = run(prompts, model_names=names, n_layer=args.layers, device=dev, shard_group=group)
    df = prepare_data_for_csv(outputs, names)
    if rank != 0:
        return df
    name = "instruct_model_comparison_results.csv" if args.survey == 1 else "instruct_model_comparison_results_survey2.csv"
    path = os.path.join(args.out, name)
    df.to_csv(path, index=False)
    log_print(f"Results saved to {path}")
    return df


if __name__ == "__main__":
    main()
