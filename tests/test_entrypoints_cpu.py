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


def test_compare_base_vs_instruct_host_logic():
    """Entry point 6: pair list, the reference's by-name prompt format (:462-463), the odds_ratio CSV schema (:90-111)."""
    from llm_interpretation_replication_b200.entrypoints import compare_base_vs_instruct as E6
    assert len(E6.model_pairs) == 9 and E6.model_pairs[0] == ("google/t5-v1_1-base", "google/flan-t5-base")
    q = 'Is a "tent" a "building"?'
    few = common.FEW_SHOT
    base_form = f"{few}Question: {q} Answer either 'Yes' or 'No', without any other text.\nAnswer:"
    bare_form = f"{few}{q} Answer either 'Yes' or 'No', without any other text."
    assert E6.format_prompt("stabilityai/stablelm-base-alpha-7b", q) == base_form
    assert E6.format_prompt("togethercomputer/RedPajama-INCITE-7B-Base", q) == base_form
    assert E6.format_prompt("bigscience/bloom-7b1", q) == base_form
    assert E6.format_prompt("tiiuae/falcon-7b", q) == bare_form              # no "base" in the name: the reference's quirk
    assert E6.format_prompt("tiiuae/falcon-7b-instruct", q) == bare_form
    out = {"tiiuae/falcon-7b": {q: {"completion": "x", "yes_prob": .2, "no_prob": .1, "odds_ratio": 2.0}},
           "tiiuae/falcon-7b-instruct": {q: {"yes_prob": float("nan"), "no_prob": float("nan"), "odds_ratio": float("nan"),
                                             "prompt": q, "completion": "N/A"}}}
    df = E6.prepare_data_for_csv(out, [("tiiuae/falcon-7b", "tiiuae/falcon-7b-instruct")])
    assert list(df.columns) == ["prompt", "model", "model_family", "base_or_instruct", "model_output", "yes_prob", "no_prob",
                                "odds_ratio"]
    assert df["base_or_instruct"].tolist() == ["base", "instruct"] and df["model_family"].tolist() == ["falcon", "falcon"]
    # the shipped table of the reference has exactly this schema
    ref = pd.read_csv(os.path.join(GOLDEN, "model_comparison_results.csv"))
    assert list(ref.columns) == list(df.columns)
    assert len(E1.MODEL_PAIRS) == 6 and [p[2] for p in E1.MODEL_PAIRS] == ["Falcon", "StableLM", "RedPajama", "BLOOM", "Pythia-Dolly", "Mistral"]
    from llm_interpretation_replication_b200 import shapes
    for pair in E1.MODEL_PAIRS:
        assert pair[0] in shapes.HUB_TO_SHAPE and pair[1] in shapes.HUB_TO_SHAPE, pair


def test_perturbation_requests_follow_the_reference_template():
    """Entry point 7: 5 scenarios x 2000 rephrasings, prompt = f"{rephrasing} {response_format}" (perturb_prompts.py:215),
    one candidate pair per scenario from target_tokens, the duplicate counts SURVEY.md 8d lists."""
    from llm_interpretation_replication_b200.entrypoints import score_perturbations as E7
    from llm_interpretation_replication_b200.tokenizer import HashTokenizer
    from llm_interpretation_replication_b200 import scorer as S
    sc = E7.load_perturbations()
    assert len(sc) == 5 and all(len(s["rephrasings"]) == 2000 for s in sc)
    assert [len(s["rephrasings"]) - len(set(s["rephrasings"])) for s in sc] == [60, 23, 28, 19, 48]
    reqs = E7.build_requests(sc)
    assert len(reqs) == 10000
    p, cands, si, reph = reqs[2000]
    assert si == 1 and p == f"{reph} {sc[1]['response_format']}" and cands == (" Ultimate", " First")
    tok = HashTokenizer(65024, eos_token_id=11)
    L = np.array([len(tok(q[0]).input_ids) for q in reqs])
    assert L.min() >= 64 and L.max() <= 191 and abs(L.sum() - 1.11e6) < 2e4          # the length histogram the bench uses
    for s in sc:
        a, b = S.candidate_ids(tok, "100q", " " + s["target_tokens"][0], " " + s["target_tokens"][1])
        assert a != b
    assert len(E7.build_requests(sc, limit=3)) == 15


def test_default_mode_per_family(tmp_path, monkeypatch):
    """load_model's default: the fastest mode that holds the parity gate for the family (and the reference's 8-bit switch)."""
    import json
    from llm_interpretation_replication_b200 import shapes
    from llm_interpretation_replication_b200.entrypoints.common import default_mode
    monkeypatch.delenv("B2S_USE_8BIT", raising=False)
    assert default_mode("tiiuae/falcon-7b-instruct") == shapes.MODE_FP16
    assert default_mode("togethercomputer/RedPajama-INCITE-7B-Instruct") == shapes.MODE_FP16
    assert default_mode("bigscience/bloomz-7b1") == shapes.MODE_FP16 and default_mode("facebook/opt-iml-1.3b") == shapes.MODE_FP16
    for name in ("mistralai/Mistral-7B-Instruct-v0.3", "meta-llama/Llama-2-7b-hf", "Qwen/Qwen-7B-Chat", "baichuan-inc/Baichuan2-7B-Chat"):
        assert default_mode(name) == shapes.MODE_BF16_SPLIT_ATTN32, name
    assert default_mode("bigscience/T0_3B") == shapes.MODE_FP16_FFN_SPLIT and default_mode("google/flan-t5-base") == shapes.MODE_FP16_FFN_SPLIT
    for mt, want in (("mistral", shapes.MODE_BF16_SPLIT_ATTN32), ("t5", shapes.MODE_FP16_FFN_SPLIT), ("gpt_neox", shapes.MODE_FP16)):
        d = tmp_path / mt
        d.mkdir()
        json.dump({"model_type": mt}, open(d / "config.json", "w"))
        assert default_mode("whatever", str(d)) == want
    monkeypatch.setenv("B2S_USE_8BIT", "1")
    assert default_mode("tiiuae/falcon-7b") == shapes.MODE_INT8_EMU and default_mode("meta-llama/Llama-2-7b-hf") == shapes.MODE_INT8_EMU
    assert default_mode("bigscience/T0_3B") == shapes.MODE_FP16_FFN_SPLIT
