"""SURVEY 8f row N3 on the CPU: the correlation / kappa oracle against the goldens produced by the reference's own
functions (tests/golden/make_corr_golden.py)."""
import json
import os

import numpy as np
import pandas as pd
import pytest

from oracle import correlation_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "corr_golden.json")))
DF = pd.read_csv(os.path.join(HERE, "golden", "corr_input.csv"))


@pytest.mark.parametrize("key", ["pearson_1000", "pearson_20", "spearman_20"])
def test_oracle_matches_reference_correlation_bootstrap(key):
    method, n_samples = key.split("_")
    r, e = O.calculate_model_correlations(DF, method, int(n_samples), 1000), GOLD[key]
    for k in ("mean_correlation", "median_correlation", "std_correlation", "min_correlation", "max_correlation", "mean_se",
              "median_se", "std_se"):
        assert abs(r[k] - e[k]) <= 1e-12, k
    for k in ("mean_ci", "median_ci", "std_ci"):
        assert np.allclose(r[k], e[k], rtol=0, atol=1e-12), k
    for k in ("bootstrap_means", "bootstrap_medians", "bootstrap_stds", "correlation_values"):
        assert len(r[k]) == len(e[k]) and np.allclose(r[k], e[k], rtol=0, atol=1e-12), k


def test_oracle_matches_reference_kappas():
    k, e = O.calculate_cohens_kappa(DF), GOLD["kappa"]
    assert np.allclose(k["kappa_scores"], e["kappa_scores"], rtol=0, atol=1e-12)
    for name in ("mean_kappa", "median_kappa", "std_kappa", "min_kappa", "max_kappa"):
        assert abs(k[name] - e[name]) <= 1e-12
    # aggregate kappa draws from the stream the spearman bootstrap left behind (reference main(): no re-seeding)
    rng = np.random.RandomState(42)
    P = DF["prompt"].nunique()
    for _ in range(1000):
        rng.choice(P, size=P, replace=True)
    a, e = O.calculate_aggregate_cohens_kappa(DF, rng=rng), GOLD["aggregate_kappa_after_spearman"]
    for name, v in e.items():
        assert abs(a[name] - v) <= 1e-12, name
