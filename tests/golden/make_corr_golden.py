#!/usr/bin/env python
"""Golden vectors for SURVEY.md 8f row N3, produced by the REFERENCE's own code.

The reference module analysis/model_comparison_graph.py imports matplotlib / seaborn (absent here) and creates
directories on a Windows drive at import time, so the three functions are taken out of its source with `ast`
(untouched) and executed with the names they use (pd, np, random, itertools, cohen_kappa_score):
  calculate_model_correlations       :207-340   (pearson and spearman, np.random.seed(42), 1000 prompt resamples)
  calculate_cohens_kappa             :495-547
  calculate_aggregate_cohens_kappa   :549-672   (continues the global NumPy stream left by the previous call)
Input: a synthetic entry-point-2 table (12 models x 60 prompts, relative_prob in [0,1], 9 missing cells, a
duplicated row) written next to the outputs.  Run in the build container (needs /root/reference):
    python tests/golden/make_corr_golden.py
Writes tests/golden/corr_input.csv and tests/golden/corr_golden.json."""
import ast
import contextlib
import io
import itertools
import json
import os
import random

import numpy as np
import pandas as pd
from sklearn.metrics import cohen_kappa_score

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/analysis/model_comparison_graph.py"
WANT = ("calculate_model_correlations", "calculate_cohens_kappa", "calculate_aggregate_cohens_kappa")


def reference_functions():
    tree = ast.parse(open(SRC).read())
    ns = dict(pd=pd, np=np, random=random, itertools=itertools, cohen_kappa_score=cohen_kappa_score)
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in WANT:
            exec(compile(ast.Module([node], []), SRC, "exec"), ns)
    return [ns[n] for n in WANT]


def make_input():
    g = np.random.RandomState(1234)
    models = [f"org{m % 4}/model-{m:02d}-instruct" for m in range(12)]
    prompts = [f'Is a "thing{p}" a "kind{p % 7}"?' for p in range(60)]
    latent = g.rand(len(prompts))
    rows = []
    for mi, m in enumerate(g.permutation(models)):
        noise = 0.15 + 0.05 * (mi % 5)
        for pi in g.permutation(len(prompts)):
            v = float(np.clip(latent[pi] + noise * g.randn(), 0.0, 1.0))
            rows.append(dict(prompt=prompts[pi], model=m, relative_prob=round(v, 6)))
    df = pd.DataFrame(rows)
    miss = g.choice(len(df), 9, replace=False)
    df.loc[miss, "relative_prob"] = np.nan                    # failed generations (NaN rows of the sweep)
    df = pd.concat([df, df.iloc[[5]].assign(relative_prob=0.25)], ignore_index=True)   # duplicate (prompt, model): pivot mean
    return df


def main():
    f_corr, f_kappa, f_agg = reference_functions()
    df = make_input()
    df.to_csv(os.path.join(HERE, "corr_input.csv"), index=False)
    df = pd.read_csv(os.path.join(HERE, "corr_input.csv"))    # what a consumer of the CSV sees
    out = {}
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        for method in ("pearson", "spearman"):
            for n_samples in (1000, 20):                      # 20 < 66 pairs: exercises random.seed(42) + random.sample
                r = f_corr(df, correlation_type=method, n_samples=n_samples, n_bootstrap=1000)
                out[f"{method}_{n_samples}"] = dict(
                    mean_correlation=r["mean_correlation"], mean_ci=list(r["mean_ci"]), mean_se=r["mean_se"],
                    median_correlation=r["median_correlation"], median_ci=list(r["median_ci"]), median_se=r["median_se"],
                    std_correlation=r["std_correlation"], std_ci=list(r["std_ci"]), std_se=r["std_se"],
                    min_correlation=r["min_correlation"], max_correlation=r["max_correlation"],
                    correlation_values=[float(v) for v in r["correlation_values"]],
                    bootstrap_means=[float(v) for v in r["bootstrap_means"]],
                    bootstrap_medians=[float(v) for v in r["bootstrap_medians"]],
                    bootstrap_stds=[float(v) for v in r["bootstrap_stds"]])
        k = f_kappa(df)
        out["kappa"] = dict(mean_kappa=k["mean_kappa"], median_kappa=k["median_kappa"], std_kappa=k["std_kappa"],
                            min_kappa=k["min_kappa"], max_kappa=k["max_kappa"], kappa_scores=[float(v) for v in k["kappa_scores"]])
        # the reference's main() calls this after the spearman bootstrap without re-seeding: reproduce that stream position
        f_corr(df, correlation_type="spearman", n_bootstrap=1000)
        a = f_agg(df, n_bootstrap=1000)
        out["aggregate_kappa_after_spearman"] = {kk: (int(v) if isinstance(v, (int, np.integer)) else float(v)) for kk, v in a.items()}
    json.dump(out, open(os.path.join(HERE, "corr_golden.json"), "w"))
    print({k: (v.get("mean_correlation", v.get("aggregate_kappa", v.get("mean_kappa")))) for k, v in out.items()})


if __name__ == "__main__":
    main()
