#!/usr/bin/env python
"""Entry point 3 -- mirrors the statistics of reference analysis/analyze_results_base_versus_instruct.py
(plots are out of scope).

Reads model_comparison_results.csv (columns prompt, model, model_family, base_or_instruct, model_output, yes_prob,
no_prob, odds_ratio), drops Mistral (:34), pairs base/instruct per family requiring all four probabilities > 0 and
recomputing relative probabilities (:38-58), pivots prompt x family (:293-298), then the prompt bootstrap and the
sign-flip permutation on one NumPy-legacy stream (:300-365) through the device kernels, and writes
overall_mae_statistics.csv (`Metric, Value` x 11 rows, :405-414).

    python -m llm_interpretation_replication_b200.entrypoints.analyze_results_base_versus_instruct CSV [--out DIR]
"""
from __future__ import annotations

import argparse
import os

import numpy as np
import pandas as pd

from .. import bootstrap


def process_model_pair(base_df, instruct_df):
    """reference :38-58"""
    merged = pd.merge(base_df, instruct_df, on="prompt", suffixes=("_base", "_instruct"))
    ok = ((merged["yes_prob_base"] > 0) & (merged["no_prob_base"] > 0) &
          (merged["yes_prob_instruct"] > 0) & (merged["no_prob_instruct"] > 0))
    merged = merged[ok].copy()
    merged["rel_prob_base"] = merged["yes_prob_base"] / (merged["yes_prob_base"] + merged["no_prob_base"])
    merged["rel_prob_instruct"] = merged["yes_prob_instruct"] / (merged["yes_prob_instruct"] + merged["no_prob_instruct"])
    return merged


def build_pivot(df):
    df = df[df["model_family"] != "mistral"]
    rows = []
    for family in df["model_family"].unique():
        fdf = df[df["model_family"] == family]
        base = fdf[fdf["base_or_instruct"] == "base"]
        inst = fdf[fdf["base_or_instruct"] == "instruct"]
        if len(base) == 0 or len(inst) == 0:
            continue
        m = process_model_pair(base, inst)
        for _, r in m.iterrows():
            rows.append({"Prompt": r["prompt"], "Model Family": family,
                         "Difference": r["rel_prob_instruct"] - r["rel_prob_base"]})
    d = pd.DataFrame(rows)
    d["Abs_Difference"] = d["Difference"].abs()
    pivot = d.pivot_table(index="Prompt", columns="Model Family", values="Difference", aggfunc="mean").dropna()
    return d, pivot


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--out", default=".")
    args = ap.parse_args(argv)
    df = pd.read_csv(args.csv)
    d, pivot = build_pivot(df)
    observed_mae = d["Abs_Difference"].mean()          # pandas means over the un-pivoted frame, as the script takes them (:327-328)
    observed_mean = d["Difference"].mean()
    r = bootstrap.prompt_bootstrap_and_signflip(pivot.values, observed_mae=observed_mae, observed_mean=observed_mean)
    stats = {"Metric": ["MAE", "MAE_CI_Lower", "MAE_CI_Upper", "MAE_P_Value", "Mean_Diff", "Mean_Diff_CI_Lower",
                        "Mean_Diff_CI_Upper", "Mean_Diff_P_Value", "N_Prompts", "N_Models", "N_Bootstrap"],
             "Value": [observed_mae, r["mae_ci_lower"], r["mae_ci_upper"], r["mae_p_value"], observed_mean,
                       r["mean_ci_lower"], r["mean_ci_upper"], r["mean_p_value"], len(pivot), pivot.shape[1], 10000]}
    os.makedirs(args.out, exist_ok=True)
    out = pd.DataFrame(stats)
    out.to_csv(os.path.join(args.out, "overall_mae_statistics.csv"), index=False)
    print(out.to_string(index=False))
    return out


if __name__ == "__main__":
    main()
