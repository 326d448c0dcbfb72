"""GPU parity of the bootstrap path: indices bit-exact vs numpy's legacy stream, statistics within 1e-6
of the reference functions' outputs (golden) -- BASELINE.json north_star."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def G():
    return json.load(open(os.path.join(GOLDEN, "bootstrap_golden.json")))


@pytest.mark.parametrize("seed,n,count", [(42, 100, 1_000_000), (42, 49, 490_000), (7, 245, 10_000), (3, 2, 5000),
                                          (9, 5, 777), (1, 1, 10), (5, 600, 60_000), (42, 128, 3000)])
def test_mt19937_choice_bit_exact(seed, n, count):
    from llm_interpretation_replication_b200.bootstrap import NumpyLegacyStream
    rs = NumpyLegacyStream(seed)
    a = rs.choice(n, count).cpu().numpy()
    np.random.seed(seed)
    b = np.random.choice(n, size=count, replace=True)
    assert np.array_equal(a, b)


def test_mt19937_stream_continues_across_calls():
    from llm_interpretation_replication_b200.bootstrap import NumpyLegacyStream
    rs = NumpyLegacyStream(42)
    parts = [rs.choice(100, 37), rs.choice(49, 1000), rs.choice(2, 5), rs.choice(300, 12345)]
    np.random.seed(42)
    for p, (n, c) in zip(parts, ((100, 37), (49, 1000), (2, 5), (300, 12345))):
        assert np.array_equal(p.cpu().numpy(), np.random.choice(n, size=c, replace=True))


def test_bootstrap_mae_difference_vs_reference(G):
    from llm_interpretation_replication_b200 import bootstrap as B
    g = G["c5_synthetic"]
    r = B.bootstrap_mae_difference(g["base"], g["instruct"])
    for k, v in g["expect"].items():
        assert abs(r[k] - v) <= 1e-6, (k, r[k], v)
    assert np.allclose(r["bootstrap_diffs"][:16], g["first_diffs"], atol=1e-12)
    assert abs(float(np.sum(r["bootstrap_diffs"])) - g["sum_diffs"]) <= 1e-8
    for fam in G["shipped_families"]:          # SURVEY.md Appendix D.1 rows (n = 49)
        r = B.bootstrap_mae_difference(fam["base"], fam["instruct"])
        for k, v in fam["expect"].items():
            assert abs(r[k] - v) <= 1e-6, (k, r[k], v)


def test_bootstrap_overall_vs_reference(G):
    from llm_interpretation_replication_b200 import bootstrap as B
    for g in (G["c5_overall_synthetic"], G["shipped_overall"]):
        r = B.bootstrap_overall_mae_difference(g["families"])
        for k, v in g["expect"].items():
            assert abs(r[k] - v) <= 1e-6, (k, r[k], v)


def test_analyze_results_and_a9_vs_reference(G):
    from llm_interpretation_replication_b200 import bootstrap as B
    g = G["analyze_results"]
    fam, overall = B.analyze_results_bootstrap({k: np.array(v) for k, v in g["family_diffs"].items()})
    for f, exp in g["expect"].items():
        for k, v in exp.items():
            assert abs(fam[f][k] - v) <= 1e-6, (f, k)
    g = G["a9_shipped"]
    r = B.prompt_bootstrap_and_signflip(np.array(g["pivot"]))
    for k, v in g["expect"].items():
        assert abs(r[k] - v) <= 1e-6, (k, r[k], v)


def test_bootstrap_properties_large_and_philox():
    """Size-independent properties at a scaled size: constant data -> every resample mean equals the constant;
    sorted output; Philox mode reproducible and statistically consistent with the MT mode."""
    from llm_interpretation_replication_b200 import bootstrap as B
    n, Bn = 100, 200_000
    boot = B.resample(np.full((2, n), 0.25), 1, Bn, None, philox_seed=123)
    assert torch.all((boot[:, 0] - 0.25).abs() < 1e-15) and torch.all((boot[:, 2] - 0.25).abs() < 1e-15)
    vals = np.random.RandomState(0).normal(size=(1, n))
    b1 = B.resample(vals, 1, Bn, None, philox_seed=5)
    b2 = B.resample(vals, 1, Bn, None, philox_seed=5)
    assert torch.equal(b1[:, [0, 2]], b2[:, [0, 2]]) and bool(torch.isnan(b1[:, 1]).all())   # no rows >= split
    s, srt = B.summary(b1, 0, 2.5, 97.5, 0.0, want_stat=True)
    x = b1[:, 0].cpu().numpy()
    assert abs(s["pct_lo"] - np.percentile(x, 2.5)) <= 1e-12 and abs(s["pct_hi"] - np.percentile(x, 97.5)) <= 1e-12
    assert abs(s["mean"] - x.mean()) <= 1e-12 and s["min"] == x.min() and s["max"] == x.max()
    assert abs(s["mean"] - vals.mean()) < 5 * vals.std() / np.sqrt(n) / np.sqrt(Bn) * 10
