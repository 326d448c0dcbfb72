"""The oracle is pinned: against the golden vectors produced by the reference's own functions
(tests/golden/make_golden.py) and against numpy's legacy RNG."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from llm_interpretation_replication_b200 import shapes, weights
from llm_interpretation_replication_b200.tokenizer import HashTokenizer
from oracle import bootstrap_oracle, decoder_ref, hf_models, scorer_oracle
from oracle.mt19937 import MT19937, choice_table_numpy


@pytest.fixture(scope="module")
def scorer_golden():
    return json.load(open(os.path.join(GOLDEN, "scorer_golden.json")))


@pytest.fixture(scope="module")
def boot_golden():
    return json.load(open(os.path.join(GOLDEN, "bootstrap_golden.json")))


def _tok(shp, row):
    tok = HashTokenizer(shp.vocab, eos_token_id=shp.eos_id).apply_pad_default()
    if row["cand_override"] is not None:
        tok.overrides = {" Yes": row["cand_override"][0], " No": row["cand_override"][1]}
    return tok


@pytest.mark.parametrize("case_idx", [0, 1, 2, 3, 4, 5, 6, 7])
def test_hf_oracle_matches_reference_function(scorer_golden, case_idx):
    case = scorer_golden["cases"][case_idx]
    shp = shapes.REGISTRY[case["shape"]]()
    W = weights.init_weights(shp, seed=case["seed"])
    model = hf_models.build_hf_model(shp, W)
    for row in case["rows"][:6]:
        tok = _tok(shp, row)
        r = scorer_oracle.get_yes_no_logprobs(model, tok, row["prompt"], variant="100q")
        g = row["ref_100q"]
        assert r["completion"] == g["completion"]
        assert abs(r["yes_prob"] - g["yes_prob"]) <= 1e-9 and abs(r["no_prob"] - g["no_prob"]) <= 1e-9
        assert abs(r["relative_prob"] - g["relative_prob"]) <= 1e-9
        r2 = scorer_oracle.get_yes_no_logprobs(model, tok, row["prompt"], variant="instruct")
        g2 = row["ref_instruct"]
        assert r2["position_found"] == g2["position_found"] and r2["yes_no_found"] == g2["yes_no_found"]
        assert abs(r2["relative_prob"] - g2["relative_prob"]) <= 1e-9


@pytest.mark.parametrize("case_idx", [0, 1, 2, 3, 4, 5, 6, 7])
def test_torch_decoder_restatement_matches_reference(scorer_golden, case_idx):
    """All 32 golden rows per family through the plain-torch restatement (k=2 and k=5 rules)."""
    case = scorer_golden["cases"][case_idx]
    shp = shapes.REGISTRY[case["shape"]]()
    W = weights.init_weights(shp, seed=case["seed"])
    for row in case["rows"]:
        tok = _tok(shp, row)
        ids = tok(row["prompt"]).input_ids
        assert len(ids) == row["n_tokens"]
        ya, na = scorer_oracle.candidate_ids(tok, "instruct")
        s2 = decoder_ref.score_ids(shp, W, ids, ya, na, topk=2, eos_id=shp.eos_id)
        g2 = row["ref_instruct"]
        assert s2["pos_found"] == g2["position_found"] and s2["found"] == g2["yes_no_found"]
        assert abs(s2["rel"] - g2["relative_prob"]) <= 1e-5
        s5 = decoder_ref.score_ids(shp, W, ids, ya, na, topk=5, eos_id=shp.eos_id)
        assert abs(s5["rel"] - row["ref_100q"]["relative_prob"]) <= 1e-5
        gen = "".join(f"<{t}>" for t in s5["gen_ids"] if t != shp.eos_id)
        assert row["ref_100q"]["completion"].startswith(gen[:100])


def test_mt19937_restatement_vs_numpy():
    for seed, n in ((42, 100), (42, 49), (7, 245), (3, 2), (9, 5), (5, 600)):
        a = MT19937(seed).choice(n, 1500)
        np.random.seed(seed)
        b = np.random.choice(n, size=1500, replace=True)
        assert (a == b).all()
    assert MT19937(42).choice(100, 8).tolist() == [51, 92, 14, 71, 60, 20, 82, 86]
    assert MT19937(42).choice(49, 8).tolist() == [38, 28, 14, 42, 7, 20, 38, 18]
    t = choice_table_numpy(42, 100, 3)
    assert t[0, :8].tolist() == [51, 92, 14, 71, 60, 20, 82, 86]


def test_bootstrap_restatement_vs_reference_golden(boot_golden):
    g = boot_golden["c5_synthetic"]
    r = bootstrap_oracle.bootstrap_mae_difference(g["base"], g["instruct"])
    for k, v in g["expect"].items():
        assert abs(r[k] - v) <= 1e-12, k
    assert np.allclose(r["bootstrap_diffs"][:16], g["first_diffs"], atol=1e-15)
    g = boot_golden["c5_overall_synthetic"]
    r = bootstrap_oracle.bootstrap_overall_mae_difference(g["families"])
    for k, v in g["expect"].items():
        assert abs(r[k] - v) <= 1e-12, k
    g = boot_golden["analyze_results"]
    fam, overall = bootstrap_oracle.analyze_results_core(g["family_diffs"])
    for f, exp in g["expect"].items():
        for k, v in exp.items():
            assert abs(fam[f][k] - v) <= 1e-12, (f, k)
    g = boot_golden["a9_shipped"]
    r = bootstrap_oracle.prompt_bootstrap_and_signflip(np.array(g["pivot"]))
    for k, v in g["expect"].items():
        assert abs(r[k] - v) <= 1e-9, k


def test_shipped_data_goldens_match_survey_appendix_d(boot_golden):
    fam = boot_golden["shipped_families"]
    assert len(fam) == 5
    assert abs(fam[0]["expect"]["base_mae"] - 0.21272931615254154) < 1e-12        # Falcon row, Appendix D.1
    assert abs(fam[0]["expect"]["ci_upper"] - 0.10571830556034265) < 1e-12
    assert abs(boot_golden["shipped_overall"]["expect"]["observed_diff"] - 0.05335786352536251) < 1e-12
    assert abs(boot_golden["a9_shipped"]["expect"]["mae"] - 0.11919688021549704) < 1e-12   # Appendix D.2
