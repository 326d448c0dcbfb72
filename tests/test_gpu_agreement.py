"""Unseeded agreement bootstraps (second half) in the Philox mode: the per-resample arithmetic equals the
NumPy / pandas restatement on the SAME index table (1e-12), the point estimates equal pandas', and the confidence intervals of
two independent streams agree within Monte-Carlo error."""
import numpy as np
import pandas as pd
import pytest
import torch

pytestmark = pytest.mark.gpu


def _llm_frame(rs, n_groups=3, n_prompts=10, n_models=6, nan_rate=0.05):
    rows, mapping = [], {}
    for g in range(1, n_groups + 1):
        base = rs.rand(n_prompts)
        for q in range(1, n_prompts + 2):                      # question 8 is the attention check the reference drops
            prompt = f"prompt_{g}_{q}"
            mapping[prompt] = f"S1_Q{g}_{q}"
            for m in range(n_models):
                if rs.rand() < nan_rate:
                    continue
                rows.append({"prompt": prompt, "model": f"m{m}", "relative_prob": float(np.clip(base[(q - 1) % n_prompts] * 0.6 + 0.4 * rs.rand(), 0, 1))})
    return pd.DataFrame(rows), mapping


def _survey_frame(rs, n_groups=3, n_q=11, per_group=40):
    cols = [f"S1_Q{g}_{q}" for g in range(1, n_groups + 1) for q in range(1, n_q + 1)]
    data = {c: [np.nan] * (n_groups * per_group) for c in cols}
    for g in range(1, n_groups + 1):
        trend = rs.rand(n_q) * 100
        for r in range((g - 1) * per_group, g * per_group):
            for q in range(1, n_q + 1):
                if q > 1 and rs.rand() < 0.12:                 # the first question decides group membership
                    continue
                data[f"S1_Q{g}_{q}"][r] = float(np.clip(0.5 * trend[q - 1] + 50 * rs.rand(), 0, 100))
    return pd.DataFrame(data), cols


def test_philox_indices_are_the_resample_kernels_stream():
    """b2s_philox_indices writes exactly the indices the Philox mode of b2s_bootstrap_resample draws."""
    from llm_interpretation_replication_b200 import agreement as A, bootstrap as Bm
    vals = np.random.RandomState(0).normal(size=(2, 37))
    idx = A.philox_indices(99, 37, 500)
    assert int(idx.min()) >= 0 and int(idx.max()) < 37
    a = Bm.resample(vals, 1, 500, None, philox_seed=99).cpu().numpy()
    b = Bm.resample(vals, 1, 500, idx.reshape(-1)).cpu().numpy()
    assert np.allclose(a, b, rtol=0, atol=1e-14)
    counts = np.bincount(idx.cpu().numpy().reshape(-1), minlength=37)
    assert counts.min() > 350 and counts.max() < 650            # 500 expected per value


def test_pearson_with_bootstrap_same_indices_and_ci():
    from llm_interpretation_replication_b200 import agreement as A
    from oracle import agreement_oracle as O
    rs = np.random.RandomState(1)
    x = rs.rand(60); y = 0.5 * x + 0.5 * rs.rand(60)
    r = A.calculate_pearson_with_bootstrap(x, y, n_bootstrap=2000, seed=5)
    idx = A.philox_indices(5, 60, 2000).cpu().numpy()
    want = O.pearson_boot_values(x, y, idx)
    assert np.allclose(np.array(r["bootstrap_distribution"]), want, rtol=0, atol=1e-12)
    from scipy.stats import pearsonr
    assert abs(r["correlation"] - pearsonr(x, y)[0]) < 1e-15
    r2 = A.calculate_pearson_with_bootstrap(x, y, n_bootstrap=2000, seed=6)          # another stream: same CI within MC error
    assert abs(r["ci_lower"] - r2["ci_lower"]) < 0.03 and abs(r["ci_upper"] - r2["ci_upper"]) < 0.03
    assert r["ci_lower"] < r["correlation"] < r["ci_upper"]


def test_llm_and_human_cross_prompt_correlations():
    from llm_interpretation_replication_b200 import agreement as A
    from oracle import agreement_oracle as O
    rs = np.random.RandomState(2)
    llm_df, mapping = _llm_frame(rs)
    res = A.calculate_llm_cross_prompt_correlations(llm_df, mapping, n_bootstrap=300, seed=11)
    # point estimates: pandas on the same pivots
    allc = []
    for g in (1, 2, 3):
        prompts = [p for p, q in mapping.items() if q.startswith(f"S1_Q{g}_") and not p.endswith("_8")]
        piv = llm_df[llm_df["prompt"].isin(prompts)].pivot_table(index="prompt", columns="model", values="relative_prob")
        cm = piv.corr().to_numpy(); iu = np.triu_indices(cm.shape[0], 1)
        allc += [c for c in cm[iu] if not np.isnan(c)]
    assert res["n_pairs"] == len(allc) and abs(res["mean_correlation"] - np.mean(allc)) < 1e-12
    # per-iteration pooled means on the same index tables
    tables = A._llm_group_tables(llm_df, mapping)
    pooled = []
    idxs = {gid: A.philox_indices(11 + 7919 * k, len(tables[gid][0]), 300).cpu().numpy() for k, gid in enumerate(sorted(tables))}
    for b in range(300):
        cs = []
        for gid in sorted(tables):
            prompts, piv = tables[gid]
            rows = piv.index.get_indexer(prompts)
            cs += O.cross_corr_mean(piv.to_numpy(dtype=np.float64), rows[idxs[gid][b]])
        pooled.append(np.mean(cs))
    assert abs(res["ci_lower"] - np.percentile(pooled, 2.5)) < 1e-10 and abs(res["ci_upper"] - np.percentile(pooled, 97.5)) < 1e-10

    df, cols = _survey_frame(rs)
    hres = A.calculate_human_cross_prompt_correlations(df, cols, n_bootstrap=200, seed=21)
    ht = A._human_group_tables(df, cols)
    pooled = []
    hidx = {gid: A.philox_indices(21 + 7919 * k, len(ht[gid][0]), 200).cpu().numpy() for k, gid in enumerate(sorted(ht))}
    for b in range(200):
        cs = []
        for gid in sorted(ht):
            cs += O.cross_corr_mean(ht[gid][1], hidx[gid][b], min_valid=5)
        pooled.append(np.mean(cs))
    assert abs(hres["ci_lower"] - np.percentile(pooled, 2.5)) < 1e-10 and abs(hres["ci_upper"] - np.percentile(pooled, 97.5)) < 1e-10
    assert hres["n_pairs"] > 1000 and hres["ci_lower"] <= hres["mean_correlation"] + 0.05

    d = A.calculate_cross_prompt_difference_ci(df, llm_df, cols, mapping, n_bootstrap=300, seed=31)
    assert d["n_bootstrap"] == 300 and d["ci_lower"] < d["mean_difference"] < d["ci_upper"]
    lo, hi = A.bootstrap_mean_ci(np.array(allc), n_bootstrap=2000, seed=3)
    assert lo < np.mean(allc) < hi


def test_model_agreement_bootstrap_membership_semantics():
    from llm_interpretation_replication_b200 import agreement as A
    from oracle import agreement_oracle as O
    rs = np.random.RandomState(4)
    qs = [f"q{i}" for i in range(50)]
    human = {q: float(rs.rand()) for q in qs}
    human["q7"] = 0.0                                           # a zero human average: its APE term is infinite and dropped
    model = {q: float(np.clip(human[q] + 0.2 * rs.randn(), 0, 1)) for q in qs if rs.rand() > 0.1}
    m = A.bootstrap_model_agreement(model, human, n_bootstrap=800, seed=9)
    idx = A.philox_indices(9, 50, 800).cpu().numpy()
    h = np.array([human[q] for q in qs]); mv = np.array([model.get(q, np.nan) for q in qs])
    want = np.array([O.agreement_metrics(h, mv, idx[b]) for b in range(800)], dtype=np.float64)
    assert m["n_bootstrap"] == 800
    for k, name in enumerate(("mae", "mse", "mape", "pearson_r")):
        assert abs(m[f"{name}_mean"] - np.nanmean(want[:, k])) < 1e-10, name
        assert abs(m[f"{name}_ci_lower"] - np.nanpercentile(want[:, k], 2.5)) < 1e-10, name
    few = {q: model[q] for q in list(model)[:6]}                # fewer than 10 matched rows in every sample
    assert A.bootstrap_model_agreement(few, human, n_bootstrap=200, seed=1) is None
    obs, lo, hi, p = A.calculate_difference_stats(rs.rand(9) + 0.3, rs.rand(9), n_bootstrap=4000, seed=2)
    assert lo < obs < hi and 0.0 <= p <= 1.0
