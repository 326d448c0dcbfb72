#!/usr/bin/env python
"""Entry point 4 -- mirrors reference survey_analysis/bootstrap_confidence_intervals.py.

Inputs (same three files, read from --dir, default cwd as the reference does, :9-17):
`consolidated_analysis_results.json` (`matching_stats.matches`), `survey_analysis_detailed.json`
(`results.by_question`), `model_comparison_results.csv`.  The reference script is not runnable as shipped (the
second file has no producer); tests/golden/make_ci_fixture.py derives both JSON files from the survey
CSV.  Steps: model table (:23-41) -> base correlations with seed 42 / 500 samples (:103-118) -> 10,000 re-seeded
iterations on the device (:126-148) -> percentile CIs (:151-172) -> difference CI (:174-184) -> per-model bootstrap
on the continuing stream (:211-230) -> bootstrap_confidence_intervals.json with the reference's schema (:277-310).
"""
from __future__ import annotations

import argparse
import json
import os
from collections import defaultdict

import numpy as np
import pandas as pd
import torch

from .. import bootstrap as B


def build_model_responses(df_models):
    mr = defaultdict(dict)
    for _, row in df_models.iterrows():
        try:
            y, n = float(row["yes_prob"]), float(row["no_prob"])
            rel = y / (y + n) if y + n > 0 else float("nan")
        except Exception:
            rel = float("nan")
        mr[row["model"]][row["prompt"]] = {"type": row["base_or_instruct"], "rel_prob": rel}
    return mr


def run(model_responses, question_mapping, by_question, n_bootstrap=10000, n_model_boot=1000):
    tables = B.build_ci_tables(model_responses, question_mapping, by_question)
    models, types = tables[0], tables[5]
    # base correlations: seed 42, 500 samples per model (:103)
    _, corrs, _ = B.simulate_individual_correlations(tables, 500, 42, 1, want_corrs=True)
    corrs = corrs.cpu().numpy()
    model_stats = {}
    for m, name in enumerate(models):
        c = corrs[m][~np.isnan(corrs[m])]
        if len(c):
            model_stats[name] = {"type": "base" if types[m] == 0 else "instruct", "mean_corr": float(np.mean(c)),
                                 "std_corr": float(np.std(c)), "n_correlations": int(len(c)), "correlations": c}
    # the re-seeded loop (:126-148): seeds 0 .. n_bootstrap-1, 100 samples per model
    means, _, stream = B.simulate_individual_correlations(tables, 100, 0, n_bootstrap, want_state=True)
    base_means = means[:, 0][~torch.isnan(means[:, 0])]
    inst_means = means[:, 1][~torch.isnan(means[:, 1])]
    base_all = np.concatenate([s["correlations"] for s in model_stats.values() if s["type"] == "base"] or [np.zeros(0)])
    inst_all = np.concatenate([s["correlations"] for s in model_stats.values() if s["type"] != "base"] or [np.zeros(0)])
    base_mean, inst_mean = float(np.mean(base_all)), float(np.mean(inst_all))
    base_ci = B.percentile_ci(base_means)
    inst_ci = B.percentile_ci(inst_means)
    k = min(base_means.numel(), inst_means.numel())
    diff_ci = B.percentile_ci(base_means[:k] - inst_means[:k])
    # per-model bootstrap on the stream the loop left behind (:211-230)
    model_cis = {}
    for name, st in model_stats.items():
        c = st["correlations"]
        n = len(c)
        idx = stream.choice(n, n_model_boot * n)
        boot = B.resample(c[None, :], 1, n_model_boot, idx, device=stream.device)
        s, _ = B.summary(boot, 0, 2.5, 97.5, 0.0)
        model_cis[name] = {"mean": st["mean_corr"], "ci_lower": s["pct_lo"], "ci_upper": s["pct_hi"], "type": st["type"]}
    return {
        "methodology": "Bootstrap confidence intervals for individual human-model correlations",
        "n_bootstrap": n_bootstrap,
        "overall_results": {
            "base": {"mean": base_mean, "ci_lower": base_ci[0], "ci_upper": base_ci[1]},
            "instruct": {"mean": inst_mean, "ci_lower": inst_ci[0], "ci_upper": inst_ci[1]},
            "difference": {"mean": base_mean - inst_mean, "ci_lower": diff_ci[0], "ci_upper": diff_ci[1],
                           "significant": bool(diff_ci[0] > 0 or diff_ci[1] < 0)}},
        "per_model_results": {m: {"type": s["type"], "mean": float(s["mean"]), "ci_lower": float(s["ci_lower"]),
                                  "ci_upper": float(s["ci_upper"])} for m, s in model_cis.items()},
    }


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--dir", default=".")
    ap.add_argument("--n-bootstrap", type=int, default=10000)
    args = ap.parse_args(argv)
    consolidated = json.load(open(os.path.join(args.dir, "consolidated_analysis_results.json")))
    detailed = json.load(open(os.path.join(args.dir, "survey_analysis_detailed.json")))
    mr = build_model_responses(pd.read_csv(os.path.join(args.dir, "model_comparison_results.csv")))
    res = run(mr, consolidated["matching_stats"]["matches"], detailed["results"]["by_question"], args.n_bootstrap)
    with open(os.path.join(args.dir, "bootstrap_confidence_intervals.json"), "w") as f:
        json.dump(res, f, indent=2)
    print(json.dumps(res["overall_results"], indent=1))
    return res


if __name__ == "__main__":
    main()
