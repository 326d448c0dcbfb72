"""NumPy restatement of reference survey_analysis/bootstrap_confidence_intervals.py (test infrastructure only).

PARITY UNPINNED: the reference script cannot run as shipped (its `survey_analysis_detailed.json` input has no
producer in the repository and the ids it looks up no longer match), so no output of the reference
exists to pin against; this restatement, executed with numpy's own legacy RNG and scipy's pearsonr exactly as the
script calls them, is the pin.  Follows :23-41 (model table), :45-52 (groups), :54-99 (simulated individuals),
:126-148 (re-seeded bootstrap), :151-156 (percentile CI), :211-230 (per-model bootstrap on the continuing stream).
"""
from __future__ import annotations

from collections import defaultdict

import numpy as np
from scipy.stats import pearsonr

GROUPS = {g: [f"Q{g}_{i}" for i in range(1, 12) if i != 8] for g in range(1, 6)}


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


def calculate_individual_correlations(model_responses, question_mapping, by_question, n_samples=100, seed=None):
    if seed is not None:
        np.random.seed(seed)
    out = defaultdict(list)
    for model, responses in model_responses.items():
        for _ in range(n_samples):
            group = np.random.randint(1, 6)
            hv, mv = [], []
            for q_id in GROUPS[group]:
                prompt = None
                for p, q in question_mapping.items():
                    if q == q_id:
                        prompt = p
                        break
                if prompt and prompt in responses and q_id in by_question:
                    mean = by_question[q_id]["mean_response"] / 100.0
                    std = by_question[q_id]["std_response"] / 100.0
                    hv.append(np.clip(np.random.normal(mean, std), 0, 1))
                    mv.append(responses[prompt]["rel_prob"])
            if len(hv) >= 8 and not any(np.isnan(mv)):
                try:
                    corr, _ = pearsonr(hv, mv)
                    if not np.isnan(corr):
                        out[model].append(corr)
                except Exception:
                    pass
    return out


def type_of(model_responses, model):
    r = model_responses[model]
    return r[list(r.keys())[0]]["type"]


def bootstrap_means(model_responses, question_mapping, by_question, n_bootstrap, n_samples=100):
    base, inst = [], []
    for i in range(n_bootstrap):
        c = calculate_individual_correlations(model_responses, question_mapping, by_question, n_samples, seed=i)
        b, s = [], []
        for model, corrs in c.items():
            if corrs:
                (b if type_of(model_responses, model) == "base" else s).extend(corrs)
        base.append(np.mean(b) if b else float("nan"))
        inst.append(np.mean(s) if s else float("nan"))
    return np.array(base), np.array(inst)


def per_model_bootstrap(corr_lists, n_iter=1000):
    """:211-230 -- uses the global stream as the loop above left it."""
    out = {}
    for model, corrs in corr_lists.items():
        means = []
        for _ in range(n_iter):
            if len(corrs) > 0:
                means.append(np.mean(np.random.choice(corrs, size=len(corrs), replace=True)))
        if means:
            out[model] = (np.percentile(means, 2.5), np.percentile(means, 97.5))
    return out
