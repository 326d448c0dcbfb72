"""TEST INFRASTRUCTURE ONLY (oracle): CPU restatement of the reference's inter-model agreement statistics
(SURVEY.md 8f row N3).  Never imported by the product package; pinned against tests/golden/corr_golden.json, which
tests/golden/make_corr_golden.py produced by running the reference's own functions.

Follows reference analysis/model_comparison_graph.py:
  calculate_model_correlations      :207-340  pivot (mean of duplicates), itertools.combinations of df['model'].unique(),
                                              random.seed(42) + random.sample when n_samples < pairs, np.random.seed(42),
                                              per resample DataFrame.corr -> np.mean / np.median / np.std of the pair values
  calculate_cohens_kappa            :495-547  sklearn cohen_kappa_score of (relative_prob > threshold) per model pair
  calculate_aggregate_cohens_kappa  :549-672  two np.random.choice calls per iteration on the global stream
DataFrame.corr is restated in NumPy: pairwise-complete observations, min_periods = 1, NaN when a variance is zero
(pandas/_libs/algos.pyx nancorr), Spearman = the same on average-tie ranks of the rows both columns hold
(nancorr_spearman)."""
import itertools
import random

import numpy as np
import pandas as pd
from scipy.stats import rankdata


def pair_corr(x: np.ndarray, y: np.ndarray, method: str) -> float:
    ok = ~(np.isnan(x) | np.isnan(y))
    x, y = x[ok], y[ok]
    if len(x) < 1:
        return np.nan
    if method == "spearman":
        x, y = rankdata(x), rankdata(y)
    x, y = x - x[0], y - y[0]          # a constant column has exactly zero variance, as in pandas' Welford pass
    dx, dy = x - x.mean(), y - y.mean()
    div = np.sqrt((dx * dx).sum() * (dy * dy).sum())
    return float((dx * dy).sum() / div) if div != 0 else np.nan


def prepare(df: pd.DataFrame, n_samples):
    """(table [P, M] in pivot order, row of every unique prompt in appearance order, column pairs, model pairs)."""
    unique_prompts = df["prompt"].unique()
    unique_models = df["model"].unique()
    pivot = df.pivot_table(index="prompt", columns="model", values="relative_prob")
    pairs = list(itertools.combinations(unique_models, 2))
    if n_samples is not None and n_samples < len(pairs):
        random.seed(42)
        pairs = random.sample(pairs, n_samples)
    rows = pivot.index.get_indexer(unique_prompts)
    cols = [(pivot.columns.get_loc(a), pivot.columns.get_loc(b)) for a, b in pairs]
    return pivot.to_numpy(dtype=np.float64), rows, np.asarray(cols, dtype=np.int32), pairs


def calculate_model_correlations(df, correlation_type="pearson", n_samples=1000, n_bootstrap=1000, confidence_level=0.95,
                                 rng=None):
    method = correlation_type.lower()
    X, rows, cols, _ = prepare(df, n_samples)
    orig = [pair_corr(X[:, i], X[:, j], method) for i, j in cols]
    orig = [c for c in orig if not np.isnan(c)]
    rng = rng if rng is not None else np.random.RandomState(42)      # np.random.seed(42), :259
    P = len(rows)
    means, medians, stds = [], [], []
    for _ in range(n_bootstrap):
        take = rows[rng.choice(P, size=P, replace=True)]
        Xb = X[take]
        c = [pair_corr(Xb[:, i], Xb[:, j], method) for i, j in cols]
        c = [v for v in c if not np.isnan(v)]
        if c:
            means.append(np.mean(c)); medians.append(np.median(c)); stds.append(np.std(c))
    alpha = 1 - confidence_level
    lo, hi = alpha / 2 * 100, (1 - alpha / 2) * 100
    return dict(mean_correlation=np.mean(orig), mean_ci=(np.percentile(means, lo), np.percentile(means, hi)), mean_se=np.std(means),
                median_correlation=np.median(orig), median_ci=(np.percentile(medians, lo), np.percentile(medians, hi)),
                median_se=np.std(medians), std_correlation=np.std(orig),
                std_ci=(np.percentile(stds, lo), np.percentile(stds, hi)), std_se=np.std(stds),
                min_correlation=np.min(orig), max_correlation=np.max(orig), correlation_values=orig,
                bootstrap_means=means, bootstrap_medians=medians, bootstrap_stds=stds, n_bootstrap=n_bootstrap,
                confidence_level=confidence_level, _rng=rng)


def cohen_kappa(a: np.ndarray, b: np.ndarray) -> float:
    """sklearn.metrics.cohen_kappa_score for two binary label vectors (labels = those present)."""
    labels = np.unique(np.concatenate([a, b]))
    n = len(labels)
    conf = np.zeros((n, n))
    for i, la in enumerate(labels):
        for j, lb in enumerate(labels):
            conf[i, j] = np.sum((a == la) & (b == lb))
    expected = np.outer(conf.sum(1), conf.sum(0)) / conf.sum()
    w = 1 - np.eye(n)
    den = (w * expected).sum()
    return float(1 - (w * conf).sum() / den) if den != 0 else float("nan")


def calculate_cohens_kappa(df, threshold=0.5):
    models = df["model"].unique()
    scores, names = [], []
    for m1, m2 in itertools.combinations(models, 2):
        d1 = df[df["model"] == m1][["prompt", "relative_prob"]]
        d2 = df[df["model"] == m2][["prompt", "relative_prob"]]
        mg = pd.merge(d1, d2, on="prompt", suffixes=("_1", "_2"))
        if len(mg) < 2:
            continue
        scores.append(cohen_kappa((mg["relative_prob_1"] > threshold).astype(int).to_numpy(),
                                  (mg["relative_prob_2"] > threshold).astype(int).to_numpy()))
        names.append((m1, m2))
    return dict(mean_kappa=np.mean(scores), median_kappa=np.median(scores), std_kappa=np.std(scores), min_kappa=np.min(scores),
                max_kappa=np.max(scores), kappa_scores=scores, model_pairs=names)


def aggregate_inputs(df, threshold=0.5):
    """(agreement rate per prompt, all binary ratings flattened, prompts used, models) -- :561-606."""
    pivot = df.pivot_table(index="prompt", columns="model", values="relative_prob")
    complete = pivot.dropna().index
    if len(complete) == 0:
        complete = pivot.dropna(thresh=2).index
    pivot = pivot.loc[complete]
    binary = (pivot > threshold).astype(int)
    rates = []
    for vals in binary.to_numpy():
        n = len(vals)
        if n < 2:
            continue
        ones = int(vals.sum())
        agree = ones * (ones - 1) // 2 + (n - ones) * (n - ones - 1) // 2
        rates.append(agree / (n * (n - 1) // 2))
    return np.asarray(rates, dtype=np.float64), binary.to_numpy().flatten().astype(np.float64), len(complete), binary.shape[1]


def calculate_aggregate_cohens_kappa(df, threshold=0.5, n_bootstrap=1000, rng=None):
    rates, allv, n_prompts, n_models = aggregate_inputs(df, threshold)
    rng = rng if rng is not None else np.random.RandomState()
    observed = np.mean(rates)
    p1 = np.mean(allv); p0 = 1 - p1
    chance = p1 * p1 + p0 * p0
    kappa = (observed - chance) / (1 - chance) if chance < 1 else 0
    samples = []
    for _ in range(n_bootstrap):
        sa = rates[rng.choice(len(rates), size=len(rates), replace=True)]
        sv = allv[rng.choice(len(allv), size=len(allv), replace=True)]
        sp1 = np.mean(sv); sp0 = 1 - sp1
        sc = sp1 * sp1 + sp0 * sp0
        if sc < 1:
            samples.append((np.mean(sa) - sc) / (1 - sc))
    lo, hi = (np.percentile(samples, 2.5), np.percentile(samples, 97.5)) if samples else (np.nan, np.nan)
    return dict(aggregate_kappa=kappa, observed_agreement=observed, chance_agreement=chance, kappa_ci_lower=lo, kappa_ci_upper=hi,
                n_prompts=n_prompts, n_models=n_models, p_class1=p1, p_class0=p0, _samples=samples)
