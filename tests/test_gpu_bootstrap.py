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


def _ci_inputs():
    import pandas as pd
    from llm_interpretation_replication_b200.entrypoints import bootstrap_confidence_intervals as E4
    fx = json.load(open(os.path.join(GOLDEN, "ci_fixture.json")))
    mr = E4.build_model_responses(pd.read_csv(os.path.join(GOLDEN, "model_comparison_results.csv")))
    return mr, fx["consolidated"]["matching_stats"]["matches"], fx["detailed"]["results"]["by_question"]


def test_ci_simulation_matches_numpy_restatement():
    """Row A11: per-seed means and the seed-42 per-sample correlations equal the NumPy/scipy restatement of
    bootstrap_confidence_intervals.py (the reference script itself cannot run as shipped: parity unpinned)."""
    from llm_interpretation_replication_b200 import bootstrap as B
    from oracle import ci_oracle as O
    mr, qm, bq = _ci_inputs()
    tables = B.build_ci_tables(mr, qm, bq)
    models = tables[0]
    means, _, _ = B.simulate_individual_correlations(tables, 100, 0, 6)
    ob, oi = O.bootstrap_means(mr, qm, bq, 6, 100)
    assert np.allclose(means[:, 0].cpu().numpy(), ob, atol=1e-12, rtol=0)
    assert np.allclose(means[:, 1].cpu().numpy(), oi, atol=1e-12, rtol=0)
    _, corrs, _ = B.simulate_individual_correlations(tables, 40, 42, 1, want_corrs=True)
    oc = O.calculate_individual_correlations(mr, qm, bq, 40, seed=42)
    corrs = corrs.cpu().numpy()
    for m, name in enumerate(models):
        got = corrs[m][~np.isnan(corrs[m])]
        want = np.array(oc.get(name, []))
        assert len(got) == len(want), name
        assert np.allclose(got, want, atol=1e-12, rtol=0), name


def test_ci_stream_continues_into_per_model_bootstrap():
    """np.random.choice after the loop continues the stream of the last seed (:211-230)."""
    from llm_interpretation_replication_b200 import bootstrap as B
    from oracle import ci_oracle as O
    mr, qm, bq = _ci_inputs()
    tables = B.build_ci_tables(mr, qm, bq)
    _, _, stream = B.simulate_individual_correlations(tables, 100, 0, 3, want_state=True)
    got = stream.choice(37, 500).cpu().numpy()
    O.bootstrap_means(mr, qm, bq, 3, 100)               # leaves numpy's global stream after seed 2's run
    want = np.random.choice(37, size=500, replace=True)
    assert np.array_equal(got, want)


def test_entry_point_4_schema(tmp_path):
    import shutil
    from llm_interpretation_replication_b200.entrypoints import bootstrap_confidence_intervals as E4
    fx = json.load(open(os.path.join(GOLDEN, "ci_fixture.json")))
    json.dump(fx["consolidated"], open(os.path.join(tmp_path, "consolidated_analysis_results.json"), "w"))
    json.dump(fx["detailed"], open(os.path.join(tmp_path, "survey_analysis_detailed.json"), "w"))
    shutil.copy(os.path.join(GOLDEN, "model_comparison_results.csv"), tmp_path)
    res = E4.main(["--dir", str(tmp_path), "--n-bootstrap", "256"])
    assert set(res) == {"methodology", "n_bootstrap", "overall_results", "per_model_results"}
    o = res["overall_results"]
    for k in ("base", "instruct", "difference"):
        assert o[k]["ci_lower"] <= o[k]["mean"] + 0.05 and o[k]["ci_upper"] >= o[k]["mean"] - 0.05
    assert len(res["per_model_results"]) >= 10          # models whose rel_probs are all finite
    assert os.path.exists(os.path.join(tmp_path, "bootstrap_confidence_intervals.json"))


@pytest.mark.parametrize("n,F,split,B", [(100, 2, 1, 10_000), (100, 6, 3, 1000), (48, 1, 1, 777), (256, 2, 1, 300), (4, 3, 0, 513),
                                         (49, 2, 1, 1000), (300, 2, 1, 500)])
def test_resample_kernels_match_numpy(n, F, split, B):
    """Index-table resampling against numpy on the same table: the staged, coalesced kernel (n % 4 == 0, n <= 256) and the
    general one (n = 49, 300), ragged last tile included."""
    from llm_interpretation_replication_b200 import bootstrap as Bm
    rs = np.random.RandomState(n + F)
    vals = rs.normal(0.1, 0.4, size=(F, n))
    idx = rs.randint(0, n, size=(B, n)).astype(np.int32)
    boot = Bm.resample(vals, split, B, torch.from_numpy(idx.reshape(-1)).cuda()).cpu().numpy()
    g = vals[:, idx]                                         # [F, B, n]
    if split > 0:
        assert np.allclose(boot[:, 0], g[:split].mean(axis=(0, 2)), rtol=0, atol=1e-13)
    else:
        assert np.isnan(boot[:, 0]).all()
    if F > split:
        assert np.allclose(boot[:, 1], g[split:].mean(axis=(0, 2)), rtol=0, atol=1e-13)
    assert np.allclose(boot[:, 2], np.abs(g).mean(axis=(0, 2)), rtol=0, atol=1e-13)


def test_mt19937_choice_large_and_state_after():
    """Several million draws in one call (many generator steps, many compaction chunks), then the stream must continue
    exactly where numpy's does -- the state handed back is the 624-word block of the last consumed word."""
    from llm_interpretation_replication_b200.bootstrap import NumpyLegacyStream
    rs = NumpyLegacyStream(123)
    np.random.seed(123)
    for n, count in ((100, 3_000_000), (7, 1), (100, 623), (65, 100_001), (2, 624 * 3)):
        got = rs.choice(n, count).cpu().numpy()
        assert np.array_equal(got, np.random.choice(n, size=count, replace=True)), (n, count)
