"""Host logic of the kept entry points (no GPU): schemas, prompt formats, pairing/pivot, checkpoint file."""
import json
import os

import numpy as np
import pandas as pd

from conftest import GOLDEN
from llm_interpretation_replication_b200.entrypoints import common
from llm_interpretation_replication_b200.entrypoints import analyze_results_base_versus_instruct as E3
from llm_interpretation_replication_b200.entrypoints import compare_instruct_models as E2
from llm_interpretation_replication_b200.entrypoints import run_base_vs_instruct_100q as E1


def test_prompt_formats_match_reference_goldens():
    g = json.load(open(os.path.join(GOLDEN, "scorer_golden.json")))
    s1, s2 = common.survey_questions()
    allq = s1 + s2
    assert len(s1) == 50 and len(s2) == 50
    rows = g["cases"][0]["rows"]
    base = [r for r in rows if r["kind"] == "base"]
    inst = [r for r in rows if r["kind"] == "instruct"]
    assert [common.format_base(q) for q in allq[:10]] == [r["prompt"] for r in base]      # built by the golden script
    assert [common.format_instruct(q) for q in allq[10:22]] == [r["prompt"] for r in inst]
    assert common.format_baichuan("Q?").startswith("<human>: Q?") and common.format_baichuan("Q?").endswith("\n<bot>:")


def test_pivot_matches_reference_script():
    """process_model_pair + pivot on the shipped CSV reproduce the matrix the reference script bootstraps."""
    g = json.load(open(os.path.join(GOLDEN, "bootstrap_golden.json")))["a9_shipped"]
    df = pd.read_csv(os.path.join(GOLDEN, "model_comparison_results.csv"))
    d, pivot = E3.build_pivot(df)
    assert list(map(str, pivot.columns)) == g["columns"]
    assert np.allclose(pivot.values, np.array(g["pivot"]), atol=1e-15)
    assert abs(d["Abs_Difference"].mean() - g["expect"]["mae"]) < 1e-15
    assert abs(d["Difference"].mean() - g["expect"]["mean_diff"]) < 1e-15


def test_csv_schema_and_checkpoint(tmp_path):
    out = {"tiiuae/falcon-7b-instruct": {"q1": {"completion": "x", "yes_prob": .1, "no_prob": .3, "relative_prob": .25},
                                         "q2": {"yes_prob": float("nan"), "no_prob": float("nan"),
                                                "relative_prob": float("nan"), "prompt": "q2", "completion": "N/A"}}}
    df = E2.prepare_data_for_csv(out, ["tiiuae/falcon-7b-instruct", "Qwen/Qwen-7B-Chat"])
    assert list(df.columns) == ["prompt", "model", "model_family", "model_output", "yes_prob", "no_prob", "relative_prob"]
    assert df["model_family"].tolist() == ["falcon", "falcon"]
    p = os.path.join(tmp_path, "ck.json")
    assert E1.load_checkpoint(p) == {"completed_models": [], "results": []}
    E1.save_checkpoint(p, {"completed_models": ["m"], "results": [{"a": 1}]})
    assert E1.load_checkpoint(p)["completed_models"] == ["m"]
    assert [m for pair in E1.MODEL_PAIRS for m in pair[:2]][0] == "tiiuae/falcon-7b"
    assert len(E2.models) == 10


def test_entry_point_5_host_pipeline_matches_reference():
    """Survey cleaning -> human averages -> per-model absolute errors: identical to the vectors the reference's
    own functions fed into its bootstrap (captured by tests/golden/make_golden.py), 1008 -> 884 respondents."""
    from llm_interpretation_replication_b200.entrypoints import analyze_base_vs_instruct_mae_100q as E5
    g = json.load(open(os.path.join(GOLDEN, "bootstrap_golden.json")))
    errs, n_resp = E5.family_errors([os.path.join(GOLDEN, "word_meaning_survey_results.csv"),
                                     os.path.join(GOLDEN, "word_meaning_survey_results_part_2.csv")],
                                    os.path.join(GOLDEN, "model_comparison_results.csv"))
    assert n_resp == 884
    assert list(errs) == ["Falcon", "StableLM", "RedPajama", "BLOOM", "Pythia-Dolly"]     # Mistral: constant values
    for (fam, e), gf in zip(errs.items(), g["shipped_families"]):
        assert np.array_equal(np.array(e["base"]), np.array(gf["base"])), fam
        assert np.array_equal(np.array(e["instruct"]), np.array(gf["instruct"])), fam
