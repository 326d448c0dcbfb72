"""NumPy / pandas restatement of the reference's UNSEEDED agreement bootstraps (test infrastructure only).

Follows survey_analysis/survey_analysis_consolidated.py:213-251 (calculate_pearson_with_bootstrap), :476-554 (the bootstrap
loop of calculate_human_cross_prompt_correlations), :616-662 (... of calculate_llm_cross_prompt_correlations) and
survey_analysis/analyze_llm_agreement_simple_bootstrap.py:89-148 (calculate_bootstrap_metrics) loop by loop, with an explicit
``numpy.random.RandomState`` in place of the reference's unseeded global one.  The reference loops cannot be pinned to a
stream (they never seed), so the GPU path is checked against these for the per-resample arithmetic (same indices -> same
numbers) and statistically for the confidence intervals.  Parity of THESE restatements with the reference is by
construction of the loops only: "parity unpinned" for the CI values themselves.
"""
from __future__ import annotations

import numpy as np
import pandas as pd
from scipy.stats import pearsonr


def pearson_boot_values(x, y, indices):
    """boot_corr for given index rows (:227-231)."""
    data = np.column_stack((x, y))
    out = []
    for idx in indices:
        bd = data[idx]
        out.append(pearsonr(bd[:, 0], bd[:, 1])[0])
    return np.array(out)


def cross_corr_mean(table: np.ndarray, idx_row, min_valid=0):
    """One bootstrap iteration for one group (:499-536 humans with min_valid=5, :635-655 models): rows idx_row of
    table [rows, cols], columns with fewer than min_valid non-NaN sampled values dropped, pandas pairwise Pearson, the
    non-NaN upper-triangle values."""
    boot = table[idx_row]
    if min_valid:
        keep = np.sum(~np.isnan(boot), axis=0) >= min_valid
        boot = boot[:, keep]
    if boot.shape[1] < 2:
        return []
    cm = pd.DataFrame(boot).corr(method="pearson").to_numpy()
    iu = np.triu_indices(cm.shape[0], 1)
    return [c for c in cm[iu] if not np.isnan(c)]


def agreement_metrics(human: np.ndarray, model: np.ndarray, idx_row, min_matched=10):
    """calculate_bootstrap_metrics (:89-148) for one sample: membership matching, MAE, MSE, MAPE %, Pearson r."""
    sampled = set(int(i) for i in idx_row)
    rows = [(human[q], model[q]) for q in range(len(human)) if q in sampled and not np.isnan(model[q])]
    if len(rows) < min_matched:
        return None
    h = np.array([r[0] for r in rows]); m = np.array([r[1] for r in rows])
    mae = np.mean(np.abs(h - m)); mse = np.mean((h - m) ** 2)
    with np.errstate(divide="ignore", invalid="ignore"):
        ape = np.abs((h - m) / h)
    ape = ape[np.isfinite(ape)]
    mape = np.mean(ape) * 100 if len(ape) else np.nan
    if len(h) > 1 and pd.Series(h).std() > 0 and pd.Series(m).std() > 0:
        r = pearsonr(h, m)[0]
    else:
        r = np.nan
    return mae, mse, mape, r
