"""NumPy restatement of the reference's seeded bootstraps (test infrastructure only).

Each function follows the cited reference lines statement by statement and uses numpy's legacy
global RNG exactly as the reference does, so index streams are identical by construction.
Pinned in tests against tests/golden/bootstrap_golden.json, which was produced by the reference's
own functions (tests/golden/make_golden.py) and against SURVEY.md Appendix D.
"""
from __future__ import annotations

import numpy as np


def bootstrap_mae_difference(base_errors, instruct_errors, n_bootstrap=10000, seed=42, confidence=0.95):
    """reference survey_analysis/analyze_base_vs_instruct_mae_100q.py:289-345"""
    np.random.seed(seed)
    base_mae = np.mean(base_errors)
    instruct_mae = np.mean(instruct_errors)
    observed_diff = instruct_mae - base_mae
    n = min(len(base_errors), len(instruct_errors))
    base = np.array(base_errors[:n])
    inst = np.array(instruct_errors[:n])
    diffs = np.empty(n_bootstrap)
    for b in range(n_bootstrap):
        idx = np.random.choice(n, size=n, replace=True)
        diffs[b] = np.mean(inst[idx]) - np.mean(base[idx])
    alpha = 1 - confidence
    ci_lower = np.percentile(diffs, (alpha / 2) * 100)
    ci_upper = np.percentile(diffs, (1 - alpha / 2) * 100)
    if observed_diff > 0:
        p = 2 * np.mean(diffs <= 0)
    else:
        p = 2 * np.mean(diffs >= 0)
    p = min(p, 1.0)
    return dict(base_mae=base_mae, instruct_mae=instruct_mae, observed_diff=observed_diff,
                ci_lower=ci_lower, ci_upper=ci_upper, p_value=p, n_questions=n, bootstrap_diffs=diffs)


def bootstrap_overall_mae_difference(all_family_errors, n_bootstrap=10000, seed=42, confidence=0.95):
    """reference survey_analysis/analyze_base_vs_instruct_mae_100q.py:348-414"""
    np.random.seed(seed)
    n = None
    for fam, e in all_family_errors.items():
        m = min(len(e["base"]), len(e["instruct"]))
        n = m if n is None else min(n, m)
    fams = list(all_family_errors.keys())
    B = np.array([all_family_errors[f]["base"][:n] for f in fams])
    I = np.array([all_family_errors[f]["instruct"][:n] for f in fams])
    obs_b, obs_i = np.mean(B), np.mean(I)
    observed_diff = obs_i - obs_b
    diffs = np.empty(n_bootstrap)
    for b in range(n_bootstrap):
        q = np.random.choice(n, size=n, replace=True)
        diffs[b] = np.mean(I[:, q]) - np.mean(B[:, q])
    alpha = 1 - confidence
    ci_lower = np.percentile(diffs, (alpha / 2) * 100)
    ci_upper = np.percentile(diffs, (1 - alpha / 2) * 100)
    if observed_diff > 0:
        p = 2 * np.mean(diffs <= 0)
    else:
        p = 2 * np.mean(diffs >= 0)
    p = min(p, 1.0)
    return dict(base_mae=obs_b, instruct_mae=obs_i, observed_diff=observed_diff, ci_lower=ci_lower,
                ci_upper=ci_upper, p_value=p, n_questions=n, n_models=len(fams), bootstrap_diffs=diffs)


def analyze_results_core(family_diffs, n_bootstrap=10000, seed=42):
    """reference analysis/run_base_vs_instruct_100q.py:614-705: ONE seed call, then per family
    (in order) n_bootstrap resamples of mean(diffs[idx]); then the same on the concatenation."""
    np.random.seed(seed)
    out = {}
    kept = {}
    for fam, diffs in family_diffs.items():
        diffs = np.asarray(diffs, dtype=np.float64)
        if len(diffs) < 10:
            continue
        kept[fam] = diffs
        obs = np.mean(diffs)
        boot = np.empty(n_bootstrap)
        for b in range(n_bootstrap):
            idx = np.random.choice(len(diffs), size=len(diffs), replace=True)
            boot[b] = np.mean(diffs[idx])
        p = 2 * np.mean(boot <= 0) if obs > 0 else 2 * np.mean(boot >= 0)
        out[fam] = dict(n_prompts=len(diffs), mean_diff=obs, mae=np.mean(np.abs(diffs)),
                        ci_lower=np.percentile(boot, 2.5), ci_upper=np.percentile(boot, 97.5),
                        p_value=min(p, 1.0))
    overall = None
    if len(kept) >= 2:
        alld = np.concatenate(list(kept.values()))
        obs = np.mean(alld)
        boot = np.empty(n_bootstrap)
        for b in range(n_bootstrap):
            idx = np.random.choice(len(alld), size=len(alld), replace=True)
            boot[b] = np.mean(alld[idx])
        p = 2 * np.mean(boot <= 0) if obs > 0 else 2 * np.mean(boot >= 0)
        overall = dict(mean_diff=obs, mae=np.mean(np.abs(alld)), ci_lower=np.percentile(boot, 2.5),
                       ci_upper=np.percentile(boot, 97.5), p_value=min(p, 1.0))
    return out, overall


def prompt_bootstrap_and_signflip(pivot, n_bootstrap=10000, seed=42):
    """reference analysis/analyze_results_base_versus_instruct.py:300-365.  ``pivot`` is the
    [n_prompts, F] matrix of instruct-minus-base rel_prob differences.  Rows are resampled; the
    statistic is taken over the flattened resample; then a sign-flip vector of n*F draws from
    choice([-1, 1]) per iteration on the SAME stream."""
    pivot = np.asarray(pivot, dtype=np.float64)
    n, F = pivot.shape
    np.random.seed(seed)
    observed_mae = np.mean(np.abs(pivot.flatten()))
    observed_mean = np.mean(pivot.flatten())
    boot_mae = np.empty(n_bootstrap)
    boot_mean = np.empty(n_bootstrap)
    for b in range(n_bootstrap):
        rows = np.random.choice(n, size=n, replace=True)
        flat = pivot[rows].flatten()
        boot_mae[b] = np.mean(np.abs(flat))
        boot_mean[b] = np.mean(flat)
    flat0 = pivot.flatten()
    perm_mae = np.empty(n_bootstrap)
    for b in range(n_bootstrap):
        signs = np.random.choice([-1, 1], size=len(flat0))
        perm_mae[b] = np.mean(np.abs(flat0 * signs))
    p_mae = np.mean(perm_mae >= observed_mae)
    if observed_mean > 0:
        p_mean = 2 * np.mean(boot_mean <= 0)
    else:
        p_mean = 2 * np.mean(boot_mean >= 0)
    return dict(mae=observed_mae, mae_ci_lower=np.percentile(boot_mae, 2.5),
                mae_ci_upper=np.percentile(boot_mae, 97.5), mae_p_value=p_mae,
                mean_diff=observed_mean, mean_ci_lower=np.percentile(boot_mean, 2.5),
                mean_ci_upper=np.percentile(boot_mean, 97.5), mean_p_value=p_mean,  # not capped (:358-361)
                n_prompts=n, n_models=F)
