"""Kept entry points end to end on the GPU (truncated random-init models), output schemas of SURVEY Appendix E."""
import json
import os

import numpy as np
import pandas as pd
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def test_entry_point_1_run_and_resume(tmp_path):
    from llm_interpretation_replication_b200.entrypoints import run_base_vs_instruct_100q as E1
    args = ["--out", str(tmp_path), "--layers", "1", "--questions", "12", "--pairs", "1"]
    df = E1.main(args)
    assert list(df.columns) == ["yes_prob", "no_prob", "relative_prob", "completion", "success", "prompt", "model",
                                "formatted_prompt", "model_family", "base_or_instruct"]
    assert len(df) == 24 and df["success"].all() and df["relative_prob"].between(0, 1).all()
    ck = json.load(open(os.path.join(tmp_path, E1.CHECKPOINT_NAME)))
    assert ck["completed_models"] == ["tiiuae/falcon-7b", "tiiuae/falcon-7b-instruct"] and len(ck["results"]) == 24
    df2 = E1.main(args)                       # everything is in the checkpoint: nothing is re-run
    assert len(df2) == 24 and np.allclose(df2["relative_prob"], df["relative_prob"])
    on_disk = pd.read_csv(os.path.join(tmp_path, E1.RESULTS_NAME))
    assert len(on_disk) == 24


def test_entry_point_2_nan_rows_for_unsupported_models(tmp_path):
    """A model outside the supported families keeps the reference's NaN rows (:502-522); the encoder-decoder models of the list
    (T0_3B, tk-instruct: reference :178-226), the Qwen-1 / Baichuan2 stand-ins (Llama block, q|k|v bias / packed W_pack; the
    Baichuan prompt format :491-492) and the other decoder-only families produce numbers."""
    from llm_interpretation_replication_b200.entrypoints import compare_instruct_models as E2
    names = ["Salesforce/xgen-7b-8k-inst", "bigscience/T0_3B", "Qwen/Qwen-7B-Chat", "baichuan-inc/Baichuan2-7B-Chat", "tiiuae/falcon-7b-instruct"]
    assert set(names[1:]) <= set(E2.models)
    out = E2.run(['Is a "tent" a "dwelling"?', 'Is "water" a "food"?'], model_names=names, n_layer=1)
    df = E2.prepare_data_for_csv(out, names)
    assert df[df.model == names[0]]["relative_prob"].isna().all()
    for name in names[1:]:
        ok = df[df.model == name]
        assert len(ok) == 2 and ok["yes_prob"].between(0, 1).all() and ok["no_prob"].between(0, 1).all(), name
        assert (ok["model_output"] != "N/A").all()
    assert df[df.model == "bigscience/T0_3B"]["model_family"].iloc[0] == "t0_3b"


def test_entry_point_3_matches_reference_output(tmp_path):
    from llm_interpretation_replication_b200.entrypoints import analyze_results_base_versus_instruct as E3
    g = json.load(open(os.path.join(GOLDEN, "bootstrap_golden.json")))["a9_shipped"]["expect"]
    out = E3.main([os.path.join(GOLDEN, "model_comparison_results.csv"), "--out", str(tmp_path)])
    v = dict(zip(out["Metric"], out["Value"]))
    for k_csv, k_g in (("MAE", "mae"), ("MAE_CI_Lower", "mae_ci_lower"), ("MAE_CI_Upper", "mae_ci_upper"),
                       ("MAE_P_Value", "mae_p_value"), ("Mean_Diff", "mean_diff"), ("Mean_Diff_CI_Lower", "mean_ci_lower"),
                       ("Mean_Diff_CI_Upper", "mean_ci_upper"), ("Mean_Diff_P_Value", "mean_p_value")):
        assert abs(v[k_csv] - g[k_g]) <= 1e-6, k_csv
    assert v["N_Prompts"] == 49 and v["N_Models"] == 3 and v["N_Bootstrap"] == 10000


def test_entry_point_5_matches_survey_appendix_d1(tmp_path):
    """base_vs_instruct_mae_results.csv on the shipped data = SURVEY.md Appendix D.1 (reference run), to 1e-6."""
    from llm_interpretation_replication_b200.entrypoints import analyze_base_vs_instruct_mae_100q as E5
    df = E5.main(["--survey", os.path.join(GOLDEN, "word_meaning_survey_results.csv"),
                  os.path.join(GOLDEN, "word_meaning_survey_results_part_2.csv"),
                  "--models", os.path.join(GOLDEN, "model_comparison_results.csv"), "--out", str(tmp_path)])
    assert list(df.columns) == ["Model_Family", "Base_MAE", "Instruct_MAE", "Difference", "CI_Lower", "CI_Upper",
                                "P_Value", "N_Questions"]
    want = {
        "Falcon": (0.21272931615254154, 0.2860028824961106, 0.07327356634356907, 0.03890079381792252, 0.10571830556034265, 0.0),
        "StableLM": (0.2463026667542609, 0.21132069200343345, -0.034981974750827455, -0.06531580056361877, -0.004964252070598784, 0.022),
        "RedPajama": (0.13668646460875206, 0.13478036259181617, -0.00190610201693589, -0.0330549058077339, 0.03000247566316429, 0.9),
        "BLOOM": (0.15951182759937632, 0.1941379214681884, 0.034626093868812075, -0.0031113518582683228, 0.07109445910330933, 0.0694),
        "Pythia-Dolly": (0.18341891794251156, 0.3791966521247064, 0.19577773418219482, 0.13710473748786867, 0.25161111176486795, 0.0),
        "OVERALL": (0.18772983861148848, 0.241087702136851, 0.05335786352536251, 0.035142340773154666, 0.07126612798098764, 0.0),
    }
    assert df["Model_Family"].tolist() == list(want)
    for _, r in df.iterrows():
        got = (r.Base_MAE, r.Instruct_MAE, r.Difference, r.CI_Lower, r.CI_Upper, r.P_Value)
        assert np.allclose(got, want[r.Model_Family], atol=1e-6), r.Model_Family
        assert r.N_Questions == 49


def test_entry_point_6_base_vs_instruct_odds_ratio_table(tmp_path):
    """compare_base_vs_instruct.py: by-name prompt formats, "base_vs" scorer variant, the T5 pair through the encoder-decoder
    scorer (NaN rows for models that cannot be run: entry point 2's test)."""
    from llm_interpretation_replication_b200.entrypoints import compare_base_vs_instruct as E6
    qs = ['Is a "tent" a "building"?', 'Is "water" a "food"?', 'Is a "drone" an "aircraft"?']
    pairs = [E6.model_pairs[0], E6.model_pairs[4]]            # T5 pair, Falcon pair
    out = E6.run(qs, pairs=pairs, n_layer=1)
    df = E6.prepare_data_for_csv(out, pairs)
    assert list(df.columns) == ["prompt", "model", "model_family", "base_or_instruct", "model_output", "yes_prob", "no_prob",
                                "odds_ratio"]
    assert len(df) == 12
    for fam, n in (("t5", 3), ("flan", 3), ("falcon", 6)):                       # "flan" is its own family under the reference's rule
        ok = df[df.model_family == fam]
        assert len(ok) == n and np.isfinite(ok["yes_prob"]).all() and (ok["odds_ratio"] > 0).all(), fam
        assert np.allclose(ok["odds_ratio"], ok["yes_prob"] / ok["no_prob"], rtol=1e-6)
        assert (ok["model_output"] != "N/A").all()
    # the encoder-decoder branch generates max_length=50 = 49 new tokens (reference :199)
    n_tok = df[df.model_family == "t5"]["model_output"].map(lambda c: c.count("<"))
    assert n_tok.max() <= 49
    # entry point 3 consumes exactly this table
    from llm_interpretation_replication_b200.entrypoints import analyze_results_base_versus_instruct as E3
    d, pivot = E3.build_pivot(df)
    assert pivot.shape == (3, 1)


def test_entry_point_7_perturbations_duplicates_identical(tmp_path):
    """All 10,000 prompts of data/perturbations.json through a 2-layer Falcon-7B-wide model: row schema, Relative_Prob as
    analyze_perturbation_results.py derives it, and identical rows for the 60/23/28/19/48 duplicate rephrasings (they sit
    in different batches next to different neighbours)."""
    from llm_interpretation_replication_b200.entrypoints import score_perturbations as E7
    df = E7.main(["--out", str(tmp_path), "--layers", "2"])
    assert len(df) == 10000
    for c in ("Model", "Original Main Part", "Response Format", "Confidence Format", "Rephrased Main Part", "Full Rephrased Prompt",
              "Model Response", "Token_1_Prob", "Token_2_Prob", "Odds_Ratio", "Relative_Prob"):
        assert c in df.columns, c
    assert np.isfinite(df["Relative_Prob"]).all() and df["Relative_Prob"].between(0, 1).all()
    assert np.allclose(df["Relative_Prob"], df["Token_1_Prob"] / (df["Token_1_Prob"] + df["Token_2_Prob"]))
    dup = df[df.duplicated(subset=["Full Rephrased Prompt"], keep=False)]
    n_dup_extra = int(df.duplicated(subset=["Full Rephrased Prompt"]).sum())
    assert n_dup_extra == 60 + 23 + 28 + 19 + 48
    for _, grp in dup.groupby("Full Rephrased Prompt"):
        assert grp["Token_1_Prob"].nunique() == 1 and grp["Token_2_Prob"].nunique() == 1 and grp["Model Response"].nunique() == 1
    assert df["Relative_Prob"].std() > 0.01
    assert os.path.exists(os.path.join(tmp_path, E7.RESULTS_NAME))


def test_entry_point_1_with_the_reference_8bit_switch(tmp_path, monkeypatch):
    """B2S_USE_8BIT=1 = the reference's USE_8BIT (run_base_vs_instruct_100q.py:84): the decoder-only models run in the LLM.int8
    emulation mode; the rows keep their schema and stay close to, but not equal to, the FP16 run's."""
    from llm_interpretation_replication_b200.entrypoints import run_base_vs_instruct_100q as E1
    args = ["--layers", "2", "--questions", "8", "--pairs", "1"]
    df16 = E1.main(args + ["--out", str(tmp_path / "fp16")])
    monkeypatch.setenv("B2S_USE_8BIT", "1")
    df8 = E1.main(args + ["--out", str(tmp_path / "int8")])
    assert list(df8.columns) == list(df16.columns) and len(df8) == len(df16) == 16 and df8["success"].all()
    d = (df8["relative_prob"] - df16["relative_prob"]).abs()
    assert 0 < d.max() < 0.2, d.max()
