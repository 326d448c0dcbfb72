"""Agreement bootstraps of the reference's survey analysis that draw from an UNSEEDED NumPy generator (SURVEY.md 8f row N3,
second half), on the GPU in the library's Philox mode.

Same function names, arguments and result dictionaries as the reference:

* ``calculate_pearson_with_bootstrap``           survey_analysis/survey_analysis_consolidated.py:213-251
* ``calculate_human_cross_prompt_correlations``  :403-554   (respondents x questions per survey group; respondents who
                                                 answered fewer than 5 of the SAMPLED questions drop out of that resample)
* ``calculate_llm_cross_prompt_correlations``    :556-674   (models x prompts per group)
* ``calculate_cross_prompt_difference_ci``       :676-806   (human mean correlation - LLM mean correlation per iteration)
* ``bootstrap_mean_ci``                          the ``np.random.randint`` mean bootstraps at :319-327 and :382-390
* ``bootstrap_model_agreement``                  survey_analysis/analyze_llm_agreement_simple_bootstrap.py:89-148,152-212
                                                 (question resampling with the script's set-membership matching)
* ``calculate_difference_stats``                 analyze_llm_agreement_simple_bootstrap.py:312-347

The reference never seeds these loops, so there is no stream to reproduce; each function takes an optional ``seed`` for the
counter-based Philox4x32-10 stream (``b2s_philox_indices`` / the Philox mode of ``b2s_bootstrap_resample``), and parity is
statistical: the point estimates are computed exactly as the reference computes them, the confidence intervals agree with
the reference's within Monte-Carlo error (tests/test_gpu_agreement.py).  What runs on the GPU is every per-resample
correlation matrix / metric (``b2s_corr_bootstrap_ex``, ``b2s_agreement_bootstrap``); grouping, pivots and percentiles of the
``n_bootstrap`` results are host NumPy / pandas, as in the reference.
"""
from __future__ import annotations

import itertools
import os
from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import _lib
from .bootstrap import resample


def _seed(seed: Optional[int]) -> int:
    return int.from_bytes(os.urandom(8), "little") >> 1 if seed is None else int(seed)


def philox_indices(seed: int, n: int, B: int, device=None) -> torch.Tensor:
    """idx [B, n] int32 on the device: the table of the Philox mode (b2s_philox_indices)."""
    lib = _lib.load()
    device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
    idx = torch.empty((B, n), dtype=torch.int32, device=device)
    with torch.cuda.device(device):
        _lib.check(lib.b2s_philox_indices(int(seed) & (2 ** 64 - 1), n, B, _lib.ptr(idx), _lib.current_stream()), "b2s_philox_indices")
    return idx


def _corr_boot(X: np.ndarray, idx: torch.Tensor, min_valid: int = 0):
    """Per resample: (mean, count) of the non-NaN pairwise Pearson correlations between the columns of X over rows idx[b]."""
    lib = _lib.load()
    P, M = X.shape
    B = idx.shape[0]
    dev = idx.device
    pairs = np.asarray(list(itertools.combinations(range(M), 2)), dtype=np.int32)
    x = torch.as_tensor(np.array(X, dtype=np.float64, order="C", copy=True)).to(dev)
    pr = torch.as_tensor(pairs).to(dev)
    corr = torch.empty((B, len(pairs)), dtype=torch.float64, device=dev)
    stats = torch.empty((B, 4), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.b2s_corr_bootstrap_ex(_lib.ptr(x), P, M, _lib.ptr(idx), B, _lib.ptr(pr), len(pairs), 0, int(min_valid),
                                             _lib.ptr(corr), _lib.ptr(stats), _lib.current_stream()), "b2s_corr_bootstrap_ex")
    st = stats.cpu().numpy()
    return st[:, 0], st[:, 3]


def _pooled_means(parts):
    """parts: [(mean [B], count [B])] per group -> mean over ALL correlations of an iteration (NaN where there are none)."""
    tot = sum(np.where(c > 0, m * c, 0.0) for m, c in parts)
    cnt = sum(c for _, c in parts)
    with np.errstate(invalid="ignore", divide="ignore"):
        return np.where(cnt > 0, tot / cnt, np.nan)


# ---- survey_analysis_consolidated.py -----------------------------------------------------------------------------------
def calculate_pearson_with_bootstrap(x, y, n_bootstrap=1000, confidence_level=0.95, seed: Optional[int] = None, device=None):
    from scipy.stats import pearsonr
    corr, p_value = pearsonr(x, y)
    data = np.column_stack((np.asarray(x, dtype=np.float64), np.asarray(y, dtype=np.float64)))
    idx = philox_indices(_seed(seed), len(data), n_bootstrap, device)          # np.random.randint(0, len(data), len(data)) per iteration
    means, _ = _corr_boot(data, idx)                                            # one pair: its correlation is the "mean"
    bootstrap_corrs = list(means)
    alpha = 1 - confidence_level
    return {"correlation": corr, "p_value": p_value,
            "ci_lower": np.percentile(bootstrap_corrs, (alpha / 2) * 100), "ci_upper": np.percentile(bootstrap_corrs, (1 - alpha / 2) * 100),
            "standard_error": np.std(bootstrap_corrs), "bootstrap_distribution": bootstrap_corrs}


def bootstrap_mean_ci(values: Sequence[float], n_bootstrap: int = 1000, seed: Optional[int] = None, device=None):
    """(ci_lower, ci_upper) of the mean of ``values`` under index resampling: the loops at survey_analysis_consolidated.py:
    319-327 and 382-390 (np.random.randint indices, np.mean of the sample, 2.5 / 97.5 percentiles)."""
    v = np.asarray(values, dtype=np.float64)
    boot = resample(v[None, :], 1, n_bootstrap, None, philox_seed=_seed(seed), device=device)[:, 0].cpu().numpy()
    return float(np.percentile(boot, 2.5)), float(np.percentile(boot, 97.5))


def _group_id(name: str):
    parts = name.split("_")
    return f"{parts[0]}_Q{parts[1][1:]}" if len(parts) >= 3 else None


def _human_group_tables(df, question_cols):
    """group id -> (questions, [questions x respondents] table in 0-1 units, NaN = unanswered) for the respondents who
    answered the group's first question (:431-447, :507-511)."""
    import pandas as pd
    groups = sorted({g for g in (_group_id(c) for c in question_cols) if g})
    out = {}
    for gid in groups:
        qs = [c for c in question_cols if c.startswith(gid + "_") and not c.endswith("_8")]
        if not qs:
            continue
        resp = df[df[qs[0]].notna()]
        if len(resp) < 2:
            continue
        tab = resp[qs].apply(pd.to_numeric, errors="coerce").to_numpy(dtype=np.float64).T / 100.0
        out[gid] = (qs, tab, resp.index)
    return out


def calculate_human_cross_prompt_correlations(df, question_cols, n_bootstrap=100, seed: Optional[int] = None, device=None):
    import pandas as pd
    all_correlations, group_results = [], {}
    tables = _human_group_tables(df, question_cols)
    kept = {}
    for gid, (qs, tab, index) in tables.items():
        keep = np.sum(~np.isnan(tab), axis=0) >= 5                     # "answered at least 5 questions" (:444-446)
        if keep.sum() < 2:
            continue
        data_df = pd.DataFrame(tab[:, keep], index=qs, columns=index[keep])
        cm = data_df.corr(method="pearson").to_numpy()
        iu = np.triu_indices(cm.shape[0], 1)
        gc = [c for c in cm[iu] if not np.isnan(c)]
        all_correlations += gc
        group_results[gid] = {"n_respondents": int(keep.sum()), "n_pairs": len(gc),
                              "mean_correlation": np.mean(gc) if gc else 0, "correlations": gc}
        kept[gid] = (qs, tab)
    base_mean = np.mean(all_correlations) if all_correlations else 0
    s = _seed(seed)
    parts = []
    for k, gid in enumerate(sorted(kept)):
        qs, tab = kept[gid]
        if len(qs) < 2:
            continue
        idx = philox_indices(s + 7919 * k, len(qs), n_bootstrap, device)
        parts.append(_corr_boot(tab, idx, min_valid=5))                # the filter is re-applied to the SAMPLED questions (:520-523)
    bootstrap_means = _pooled_means(parts) if parts else np.array([])
    bootstrap_means = bootstrap_means[~np.isnan(bootstrap_means)]
    if len(bootstrap_means):
        ci_lower, ci_upper = np.percentile(bootstrap_means, 2.5), np.percentile(bootstrap_means, 97.5)
    else:
        ci_lower = ci_upper = base_mean
    return {"group_results": group_results, "pairwise_correlations": all_correlations, "mean_correlation": base_mean,
            "std_correlation": np.std(all_correlations) if all_correlations else 0, "n_pairs": len(all_correlations),
            "ci_lower": ci_lower, "ci_upper": ci_upper}


def _llm_group_tables(llm_df, human_question_mapping):
    prompt_to_group = {}
    for prompt, survey_q in human_question_mapping.items():
        if survey_q:
            g = _group_id(survey_q)
            if g:
                prompt_to_group[prompt] = g
    out = {}
    for gid in sorted(set(prompt_to_group.values())):
        prompts = [p for p, g in prompt_to_group.items() if g == gid and not p.endswith("_8")]
        if len(prompts) < 2:
            continue
        group_data = llm_df[llm_df["prompt"].isin(prompts)]
        pivot_df = group_data.pivot_table(index="prompt", columns="model", values="relative_prob")
        if len(pivot_df) < 2:
            continue
        out[gid] = (prompts, pivot_df)
    return out


def calculate_llm_cross_prompt_correlations(llm_df, human_question_mapping, n_bootstrap=100, seed: Optional[int] = None, device=None):
    all_correlations, group_results = [], {}
    tables = _llm_group_tables(llm_df, human_question_mapping)
    for gid, (prompts, pivot_df) in tables.items():
        cm = pivot_df.corr(method="pearson")
        models = list(cm.columns)
        iu = np.triu_indices(len(models), 1)
        gc = [c for c in cm.to_numpy()[iu] if not np.isnan(c)]
        all_correlations += gc
        group_results[gid] = {"n_prompts": len(prompts), "n_models": len(models), "n_pairs": len(gc),
                              "mean_correlation": np.mean(gc) if gc else 0, "correlations": gc}
    base_mean = np.mean(all_correlations) if all_correlations else 0
    s = _seed(seed)
    parts = []
    for k, gid in enumerate(sorted(tables)):
        prompts, pivot_df = tables[gid]
        rows = pivot_df.index.get_indexer(prompts)                     # pivot_df.loc[sampled_prompts] (:645): every prompt must be a row
        if (rows < 0).any():
            raise KeyError(f"group {gid}: {int((rows < 0).sum())} prompts have no relative_prob for any model")
        draws = philox_indices(s + 7919 * k, len(prompts), n_bootstrap, device)
        idx = torch.as_tensor(rows.astype(np.int32)).to(draws.device)[draws.long()].contiguous()
        parts.append(_corr_boot(pivot_df.to_numpy(dtype=np.float64), idx))
    bootstrap_means = _pooled_means(parts) if parts else np.array([])
    bootstrap_means = bootstrap_means[~np.isnan(bootstrap_means)]
    if len(bootstrap_means):
        ci_lower, ci_upper = np.percentile(bootstrap_means, 2.5), np.percentile(bootstrap_means, 97.5)
    else:
        ci_lower = ci_upper = base_mean
    return {"group_results": group_results, "pairwise_correlations": all_correlations, "mean_correlation": base_mean,
            "std_correlation": np.std(all_correlations) if all_correlations else 0, "n_pairs": len(all_correlations),
            "ci_lower": ci_lower, "ci_upper": ci_upper}


def calculate_cross_prompt_difference_ci(df, llm_df, question_cols, human_question_mapping, n_bootstrap=1000,
                                         seed: Optional[int] = None, device=None):
    human = _human_group_tables(df, question_cols)
    llm = _llm_group_tables(llm_df, human_question_mapping)
    groups_human = {g for g in (_group_id(c) for c in question_cols) if g}
    groups_llm = {g for g in (_group_id(q) for q in human_question_mapping.values() if q) if g}
    groups = sorted(groups_human & groups_llm)                          # :697-700
    s = _seed(seed)
    h_parts, l_parts = [], []
    for k, gid in enumerate(groups):
        n_questions = len([c for c in question_cols if c.startswith(gid + "_") and not c.endswith("_8")])
        if n_questions < 2:                                              # :711-712 skips the group on both sides
            continue
        if gid in human:
            qs, tab, _ = human[gid]
            h_parts.append(_corr_boot(tab, philox_indices(s + 7919 * k, len(qs), n_bootstrap, device), min_valid=5))
        if gid in llm:
            prompts, pivot_df = llm[gid]
            rows = pivot_df.index.get_indexer(prompts)
            if (rows < 0).any():
                raise KeyError(f"group {gid}: prompts without any relative_prob")
            draws = philox_indices(s + 7919 * k + 104729, len(prompts), n_bootstrap, device)
            idx = torch.as_tensor(rows.astype(np.int32)).to(draws.device)[draws.long()].contiguous()
            l_parts.append(_corr_boot(pivot_df.to_numpy(dtype=np.float64), idx))
    if not h_parts or not l_parts:
        return {"mean_difference": None, "ci_lower": None, "ci_upper": None, "n_bootstrap": 0}
    diff = _pooled_means(h_parts) - _pooled_means(l_parts)              # iterations where either side is empty are skipped (:787)
    diff = diff[~np.isnan(diff)]
    if len(diff) == 0:
        return {"mean_difference": None, "ci_lower": None, "ci_upper": None, "n_bootstrap": 0}
    return {"mean_difference": np.mean(diff), "ci_lower": np.percentile(diff, 2.5), "ci_upper": np.percentile(diff, 97.5),
            "n_bootstrap": len(diff)}


# ---- analyze_llm_agreement_simple_bootstrap.py -----------------------------------------------------------------------------
def bootstrap_model_agreement(model_probs: Dict[str, float], human_avgs: Dict[str, float], n_bootstrap=1000,
                              confidence_level=0.95, seed: Optional[int] = None, device=None):
    """The per-model loop body of analyze_llm_agreement_simple_bootstrap.py:169-212 with calculate_bootstrap_metrics (:89-148):
    each iteration samples len(human_avgs) question indices with replacement; a model row is matched when its prompt is AMONG
    the sampled questions (membership, so a question counts once however often it was drawn); at least 10 matched rows are
    needed; MAE, MSE, MAPE (finite terms only, in percent) and Pearson r of (human average, model probability).
    ``model_probs`` maps prompt -> relative probability (NaN rows are skipped, as ``pd.notna`` does at :118).  Returns the
    metrics dict of :189-210 (``*_mean``, ``*_ci_lower``, ``*_ci_upper``, ``*_std``, ``n_bootstrap``), or None with fewer
    than 100 usable iterations (:182-184)."""
    lib = _lib.load()
    questions = list(human_avgs.keys())
    n = len(questions)
    human = np.asarray([human_avgs[q] for q in questions], dtype=np.float64)
    model = np.asarray([model_probs.get(q, np.nan) for q in questions], dtype=np.float64)
    idx = philox_indices(_seed(seed), n, n_bootstrap, device)
    dev = idx.device
    out = torch.empty((n_bootstrap, 4), dtype=torch.float64, device=dev)
    h, m = torch.as_tensor(human).to(dev), torch.as_tensor(model).to(dev)
    with torch.cuda.device(dev):
        _lib.check(lib.b2s_agreement_bootstrap(_lib.ptr(h), _lib.ptr(m), n, _lib.ptr(idx), n_bootstrap, 10, _lib.ptr(out),
                                               _lib.current_stream()), "b2s_agreement_bootstrap")
    res = out.cpu().numpy()
    res = res[~np.isnan(res[:, 0])]                                    # iterations with fewer than 10 matched rows return None
    if len(res) < 100:
        return None
    alpha = 1 - confidence_level
    metrics = {"n_bootstrap": len(res)}
    for k, name in enumerate(("mae", "mse", "mape", "pearson_r")):
        vals = res[:, k][~np.isnan(res[:, k])]
        if len(vals):
            metrics[f"{name}_mean"] = np.mean(vals)
            metrics[f"{name}_ci_lower"] = np.percentile(vals, (alpha / 2) * 100)
            metrics[f"{name}_ci_upper"] = np.percentile(vals, (1 - alpha / 2) * 100)
            metrics[f"{name}_std"] = np.std(vals)
        else:
            metrics[f"{name}_mean"] = metrics[f"{name}_ci_lower"] = metrics[f"{name}_ci_upper"] = metrics[f"{name}_std"] = np.nan
    return metrics


def calculate_difference_stats(group1_values, group2_values, n_bootstrap=10000, seed: Optional[int] = None, device=None):
    """analyze_llm_agreement_simple_bootstrap.py:312-347: observed difference of means, bootstrap CI of the difference (each
    group resampled with replacement) and a permutation p-value (random relabelling of the pooled values)."""
    g1, g2 = np.asarray(group1_values, dtype=np.float64), np.asarray(group2_values, dtype=np.float64)
    n1, n2 = len(g1), len(g2)
    observed_diff = np.mean(g1) - np.mean(g2)
    s = _seed(seed)
    m1 = resample(g1[None, :], 1, n_bootstrap, None, philox_seed=s, device=device)[:, 0].cpu().numpy()
    m2 = resample(g2[None, :], 1, n_bootstrap, None, philox_seed=s + 1, device=device)[:, 0].cpu().numpy()
    diffs = m1 - m2
    ci_lower, ci_upper = np.percentile(diffs, 2.5), np.percentile(diffs, 97.5)
    # permutation test: a random relabelling = a random permutation of the pooled values; n1 + n2 is a dozen models, so the
    # n_bootstrap permutations are drawn on the host (this is not a resampling-with-replacement loop the kernels cover)
    pooled = np.concatenate([g1, g2])
    rs = np.random.Generator(np.random.Philox(key=s))
    perm = rs.permuted(np.tile(pooled, (n_bootstrap, 1)), axis=1)
    null = perm[:, :n1].mean(axis=1) - perm[:, n1:].mean(axis=1)
    p_value = np.mean(np.abs(null) >= np.abs(observed_diff))
    return observed_diff, ci_lower, ci_upper, p_value
