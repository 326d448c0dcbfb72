"""Inter-model agreement statistics of the reference's comparison script on the GPU.

Same function names, arguments and result dictionaries as reference ``analysis/model_comparison_graph.py``:

* ``calculate_model_correlations``      :207-340.  The 1000 ``bootstrap_pivot.corr(method)`` calls -- one pandas
  correlation matrix per prompt resample -- become ONE ``b2s_corr_bootstrap`` launch sequence: the index stream of
  ``np.random.seed(42)`` + ``np.random.choice(n_prompts, n_prompts)`` is generated on the device bit for bit
  (``b2s_mt19937_choice``), every (resample, model pair) correlation is computed there (pairwise-complete Pearson, or
  Spearman on average-tie ranks), and so are ``np.mean / np.median / np.std`` of each resample's pair values.  The
  percentiles and standard errors of the resulting three vectors of length ``n_bootstrap`` are NumPy on the host, as
  is everything the reference does once (pivot table, ``random.sample`` of the pairs, the original matrix).
* ``calculate_cohens_kappa``            :495-547.  One pass over the model pairs, no resampling: host NumPy.
* ``calculate_aggregate_cohens_kappa``  :549-672.  Two ``np.random.choice`` draws per iteration on the GLOBAL NumPy
  stream, which in the reference's ``main`` is whatever the last seeded correlation bootstrap left behind; here the
  stream object is explicit (``stream=``; by default the one the last ``calculate_model_correlations`` call left),
  the draws are generated on the device in the same order and the gathered means use ``b2s_bootstrap_resample``.
"""
from __future__ import annotations

import itertools
import random
from typing import Optional

import numpy as np
import torch

from . import _lib
from .bootstrap import NumpyLegacyStream, resample

_last_stream: Optional[NumpyLegacyStream] = None     # stands in for NumPy's global RandomState across calls


def corr_bootstrap(X: np.ndarray, idx: torch.Tensor, pairs: np.ndarray, method: str, device=None):
    """(corr [B, n_pairs], stats [B, 4] = mean, median, std, count) for the resamples ``idx`` [B, P] of table ``X``."""
    lib = _lib.load()
    device = torch.device(device if device is not None else idx.device)
    P, M = X.shape
    B = idx.numel() // P
    x = torch.as_tensor(np.array(X, dtype=np.float64, order="C", copy=True)).to(device)
    pr = torch.as_tensor(np.ascontiguousarray(pairs, dtype=np.int32)).to(device)
    corr = torch.empty((B, pr.shape[0]), dtype=torch.float64, device=device)
    stats = torch.empty((B, 4), dtype=torch.float64, device=device)
    code = {"pearson": 0, "spearman": 1}[method]
    with torch.cuda.device(device):
        _lib.check(lib.b2s_corr_bootstrap(_lib.ptr(x), P, M, _lib.ptr(idx), B, _lib.ptr(pr), pr.shape[0], code, _lib.ptr(corr),
                                          _lib.ptr(stats), _lib.current_stream()), "b2s_corr_bootstrap")
    return corr, stats


def calculate_model_correlations(df, correlation_type="pearson", n_samples=1000, n_bootstrap=1000, confidence_level=0.95,
                                 device=None):
    global _last_stream
    method = correlation_type.lower()
    method = method if method == "pearson" else "spearman"        # the reference's else-branch (:272-275)
    unique_prompts = df["prompt"].unique()
    unique_models = df["model"].unique()
    pivot_df = df.pivot_table(index="prompt", columns="model", values="relative_prob")
    model_pairs = list(itertools.combinations(unique_models, 2))
    if n_samples is not None and n_samples < len(model_pairs):
        random.seed(42)
        model_pairs = random.sample(model_pairs, n_samples)

    original_corr_matrix = pivot_df.corr(method=method)
    original = [original_corr_matrix.loc[a, b] for a, b in model_pairs]
    original = [c for c in original if not np.isnan(c)]

    rows = pivot_df.index.get_indexer(unique_prompts)
    if (rows < 0).any():      # a prompt without any value is dropped by the pivot; the reference's .loc raises here too
        raise KeyError(f"{int((rows < 0).sum())} prompts have no relative_prob for any model")
    cols = np.asarray([(pivot_df.columns.get_loc(a), pivot_df.columns.get_loc(b)) for a, b in model_pairs], dtype=np.int32)
    n_prompts = len(unique_prompts)

    stream = NumpyLegacyStream(42, device=device)                                        # np.random.seed(42), :259
    draws = stream.choice(n_prompts, n_bootstrap * n_prompts)                            # n_bootstrap x np.random.choice
    idx = torch.as_tensor(rows.astype(np.int32)).to(draws.device)[draws.long()].contiguous()   # unique_prompts[...] -> pivot rows
    _, stats = corr_bootstrap(pivot_df.to_numpy(dtype=np.float64), idx, cols, method, device=draws.device)
    _last_stream = stream
    st = stats.cpu().numpy()
    st = st[st[:, 3] > 0]                                                                # `if bootstrap_correlations:`
    means, medians, stds = st[:, 0], st[:, 1], st[:, 2]

    alpha = 1 - confidence_level
    lo, hi = (alpha / 2) * 100, (1 - alpha / 2) * 100
    return {
        "mean_correlation": np.mean(original), "mean_ci": (np.percentile(means, lo), np.percentile(means, hi)),
        "mean_se": np.std(means),
        "median_correlation": np.median(original), "median_ci": (np.percentile(medians, lo), np.percentile(medians, hi)),
        "median_se": np.std(medians),
        "std_correlation": np.std(original), "std_ci": (np.percentile(stds, lo), np.percentile(stds, hi)),
        "std_se": np.std(stds),
        "min_correlation": np.min(original), "max_correlation": np.max(original),
        "correlation_matrix": original_corr_matrix, "correlation_values": original,
        "bootstrap_means": list(means), "bootstrap_medians": list(medians), "bootstrap_stds": list(stds),
        "n_bootstrap": n_bootstrap, "confidence_level": confidence_level,
    }


def _cohen_kappa(a: np.ndarray, b: np.ndarray) -> float:
    """sklearn.metrics.cohen_kappa_score (unweighted) for two label vectors."""
    labels = np.unique(np.concatenate([a, b]))
    conf = np.array([[np.sum((a == la) & (b == lb)) for lb in labels] for la in labels], dtype=np.float64)
    expected = np.outer(conf.sum(1), conf.sum(0)) / conf.sum()
    w = 1 - np.eye(len(labels))
    den = (w * expected).sum()
    return float(1 - (w * conf).sum() / den) if den != 0 else float("nan")


def calculate_cohens_kappa(df, threshold=0.5):
    import pandas as pd
    unique_models = df["model"].unique()
    kappa_scores, model_pair_names = [], []
    for model1, model2 in itertools.combinations(unique_models, 2):
        data1 = df[df["model"] == model1][["prompt", "relative_prob"]]
        data2 = df[df["model"] == model2][["prompt", "relative_prob"]]
        merged = pd.merge(data1, data2, on="prompt", suffixes=("_1", "_2"))
        if len(merged) < 2:
            continue
        kappa_scores.append(_cohen_kappa((merged["relative_prob_1"] > threshold).astype(int).to_numpy(),
                                         (merged["relative_prob_2"] > threshold).astype(int).to_numpy()))
        model_pair_names.append((model1, model2))
    return {"mean_kappa": np.mean(kappa_scores), "median_kappa": np.median(kappa_scores), "std_kappa": np.std(kappa_scores),
            "min_kappa": np.min(kappa_scores), "max_kappa": np.max(kappa_scores), "kappa_scores": kappa_scores,
            "model_pairs": model_pair_names}


def calculate_aggregate_cohens_kappa(df, threshold=0.5, n_bootstrap=1000, stream: Optional[NumpyLegacyStream] = None,
                                     device=None):
    pivot_df = df.pivot_table(index="prompt", columns="model", values="relative_prob")
    complete_prompts = pivot_df.dropna().index
    if len(complete_prompts) == 0:
        complete_prompts = pivot_df.dropna(thresh=2).index
    pivot_df = pivot_df.loc[complete_prompts]
    binary = (pivot_df > threshold).astype(int).to_numpy()
    n_models = binary.shape[1]
    agreement_rates = []
    if n_models >= 2:
        ones = binary.sum(axis=1)
        agree = ones * (ones - 1) // 2 + (n_models - ones) * (n_models - ones - 1) // 2
        agreement_rates = list(agree / (n_models * (n_models - 1) // 2))
    rates = np.asarray(agreement_rates, dtype=np.float64)
    observed_agreement = np.mean(rates)
    all_values = binary.flatten().astype(np.float64)
    p1 = np.mean(all_values)
    p0 = 1 - p1
    chance_agreement = p1 * p1 + p0 * p0
    kappa = (observed_agreement - chance_agreement) / (1 - chance_agreement) if chance_agreement < 1 else 0

    if stream is None:
        stream = _last_stream
    if stream is None:        # the reference is unseeded when nothing seeded the global stream before: any seed will do
        stream = NumpyLegacyStream(int.from_bytes(__import__("os").urandom(4), "little"), device=device)
    nA, nV = len(rates), len(all_values)
    dev = stream.device
    idx_a = torch.empty((n_bootstrap, nA), dtype=torch.int32, device=dev)
    idx_v = torch.empty((n_bootstrap, nV), dtype=torch.int32, device=dev)
    for i in range(n_bootstrap):                         # draw order of :616-622: agreements first, then all values
        stream.choice_into(nA, idx_a[i])
        stream.choice_into(nV, idx_v[i])
    mean_a = resample(rates[None, :], 1, n_bootstrap, idx=idx_a.reshape(-1), device=dev)[:, 0].cpu().numpy()
    mean_v = resample(all_values[None, :], 1, n_bootstrap, idx=idx_v.reshape(-1), device=dev)[:, 0].cpu().numpy()
    sample_chance = mean_v * mean_v + (1 - mean_v) * (1 - mean_v)
    keep = sample_chance < 1
    kappa_samples = (mean_a[keep] - sample_chance[keep]) / (1 - sample_chance[keep])
    if len(kappa_samples) > 0:
        lo, hi = np.percentile(kappa_samples, 2.5), np.percentile(kappa_samples, 97.5)
    else:
        lo = hi = np.nan
    return {"aggregate_kappa": kappa, "observed_agreement": observed_agreement, "chance_agreement": chance_agreement,
            "kappa_ci_lower": lo, "kappa_ci_upper": hi, "n_prompts": len(complete_prompts), "n_models": n_models,
            "p_class1": p1, "p_class0": p0}
