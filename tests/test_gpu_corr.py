"""SURVEY 8f row N3 on the GPU: b2s_corr_bootstrap and the host functions built on it against the goldens produced by
the reference's own functions, and against the NumPy oracle on seeded tables with missing cells, ties and constant
columns."""
import json
import os

import numpy as np
import pandas as pd
import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "corr_golden.json")))
DF = pd.read_csv(os.path.join(HERE, "golden", "corr_input.csv"))
TOL = 1e-9


@pytest.mark.parametrize("key", ["pearson_1000", "pearson_20", "spearman_1000", "spearman_20"])
def test_model_correlations_vs_reference_golden(key):
    from llm_interpretation_replication_b200 import correlations as Cn
    method, n_samples = key.split("_")
    r, e = Cn.calculate_model_correlations(DF, method, int(n_samples), 1000), GOLD[key]
    for k in ("mean_correlation", "median_correlation", "std_correlation", "min_correlation", "max_correlation", "mean_se",
              "median_se", "std_se"):
        assert abs(r[k] - e[k]) <= TOL, k
    for k in ("mean_ci", "median_ci", "std_ci"):
        assert np.allclose(r[k], e[k], rtol=0, atol=TOL), k
    for k in ("bootstrap_means", "bootstrap_medians", "bootstrap_stds", "correlation_values"):
        assert len(r[k]) == len(e[k]) and np.allclose(r[k], e[k], rtol=0, atol=TOL), k
    assert r["correlation_matrix"].shape == (12, 12) and r["n_bootstrap"] == 1000


def test_kappas_vs_reference_golden():
    """Pairwise kappa, then the aggregate kappa on the stream position the spearman bootstrap leaves (reference main())."""
    from llm_interpretation_replication_b200 import correlations as Cn
    k, e = Cn.calculate_cohens_kappa(DF), GOLD["kappa"]
    assert np.allclose(k["kappa_scores"], e["kappa_scores"], rtol=0, atol=1e-12)
    Cn.calculate_model_correlations(DF, "spearman", 1000, 1000)
    a, e = Cn.calculate_aggregate_cohens_kappa(DF, n_bootstrap=1000), GOLD["aggregate_kappa_after_spearman"]
    for name, v in e.items():
        assert abs(a[name] - v) <= TOL, name


@pytest.mark.parametrize("method", ["pearson", "spearman"])
@pytest.mark.parametrize("P,M,B", [(1, 3, 4), (2, 2, 5), (37, 5, 16), (300, 18, 8), (2048, 4, 3)])
def test_corr_kernel_vs_numpy_oracle(method, P, M, B):
    """Every (resample, pair) value, NaN pattern included: missing cells, heavy ties, a constant column, an empty pair."""
    from llm_interpretation_replication_b200 import correlations as Cn
    from oracle import correlation_oracle as O
    g = np.random.RandomState(P * 131 + M)
    X = np.round(g.rand(P, M), 1 if P > 30 else 3)                 # one decimal -> many ties
    X[g.rand(P, M) < 0.1] = np.nan
    if M >= 4:
        X[:, 1] = 0.5                                               # constant column -> NaN correlations
        X[:, 2] = np.nan if P < 40 else X[:, 2]                     # column nobody answered
    idx = g.randint(0, P, size=(B, P)).astype(np.int32)
    pairs = np.array([(i, j) for i in range(M) for j in range(i + 1, M)], dtype=np.int32)
    corr, stats = Cn.corr_bootstrap(X, torch.as_tensor(idx).cuda().reshape(-1), pairs, method)
    corr, stats = corr.cpu().numpy(), stats.cpu().numpy()
    for b in range(B):
        Xb = X[idx[b]]
        want = np.array([O.pair_corr(Xb[:, i], Xb[:, j], method) for i, j in pairs])
        assert np.array_equal(np.isnan(want), np.isnan(corr[b])), (b, want, corr[b])
        ok = ~np.isnan(want)
        assert np.allclose(corr[b][ok], want[ok], rtol=0, atol=TOL)
        if ok.any():
            assert abs(stats[b, 0] - np.mean(want[ok])) <= TOL and abs(stats[b, 1] - np.median(want[ok])) <= TOL
            assert abs(stats[b, 2] - np.std(want[ok])) <= TOL and stats[b, 3] == ok.sum()
        else:
            assert np.isnan(stats[b, :3]).all() and stats[b, 3] == 0


def test_corr_bootstrap_rejects_bad_arguments():
    from llm_interpretation_replication_b200 import _lib, correlations as Cn
    X = np.random.rand(3000, 3)
    idx = torch.zeros(3000, dtype=torch.int32, device="cuda")
    pairs = np.array([(0, 1)], dtype=np.int32)
    Cn.corr_bootstrap(X, idx, pairs, "pearson")                     # pearson has no row limit
    with pytest.raises(_lib.B2SError, match="2048"):
        Cn.corr_bootstrap(X, idx, pairs, "spearman")
