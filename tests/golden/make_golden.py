#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by running the REFERENCE's own code.

Run in the build container only (needs /root/reference, read-only):
    python tests/golden/make_golden.py
Writes scorer_golden.json and bootstrap_golden.json next to this file.  Nothing here is read
at test time except the JSON outputs; /root/reference does not exist on the GPU box.

What is executed from the reference:
  * analysis/run_base_vs_instruct_100q.py           -> get_yes_no_logprobs (imported module object),
                                                        analyze_results(df)
  * analysis/compare_instruct_models.py:171-293      -> get_yes_no_logprobs (k=2 variant; extracted with
                                                        ``ast`` because importing the script runs the sweep)
  * survey_analysis/analyze_base_vs_instruct_mae_100q.py -> bootstrap_mae_difference,
        bootstrap_overall_mae_difference and the survey-cleaning pipeline on the shipped data
  * analysis/analyze_results_base_versus_instruct.py -> whole script through runpy with plotting stubbed
They are driven with random-init HF models built from llm_interpretation_replication_b200.weights
(seeded, so tests can rebuild the identical weights) and the offline HashTokenizer.
"""
import ast
import contextlib
import io
import json
import os
import runpy
import sys
import tempfile
import types
from unittest.mock import MagicMock

import numpy as np
import pandas as pd
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"

from llm_interpretation_replication_b200 import shapes, weights  # noqa: E402
from llm_interpretation_replication_b200.tokenizer import HashTokenizer  # noqa: E402
from oracle import hf_models  # noqa: E402

FEW_SHOT = '''Question: Is "soup" a "beverage"? Answer either 'Yes' or 'No', without any other text.
Answer: No.

Question: Is a "tweet" a "publication"? Answer either 'Yes' or 'No', without any other text.
Answer: Yes.

'''


def import_reference_100q(scratch):
    os.chdir(scratch)
    sys.path.insert(0, os.path.join(REF, "analysis"))
    with contextlib.redirect_stdout(io.StringIO()):
        import run_base_vs_instruct_100q as R
    return R


def extract_k2_scorer():
    src = open(os.path.join(REF, "analysis", "compare_instruct_models.py")).read()
    tree = ast.parse(src)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "get_yes_no_logprobs"][0]
    mod = ast.Module(body=[fn], type_ignores=[])
    ns = {"torch": torch, "np": np, "MAX_LOOK_AHEAD": 10}
    # MAX_LOOK_AHEAD is a module constant in the script (compare_instruct_models.py:173 region)
    for n in tree.body:
        if isinstance(n, ast.Assign) and getattr(n.targets[0], "id", "") == "MAX_LOOK_AHEAD":
            ns["MAX_LOOK_AHEAD"] = ast.literal_eval(n.value)
    exec(compile(mod, "compare_instruct_models.py", "exec"), ns)
    return ns["get_yes_no_logprobs"]


SPECS = [
    ("tiny-falcon", shapes.tiny_falcon(), 101),
    ("tiny-neox-par", shapes.tiny_neox_parallel(), 102),
    ("tiny-neox-seq", shapes.tiny_neox_sequential(), 103),
    # Llama family (Mistral-7B-Instruct of the instruct sweep, compare_instruct_models.py:157): added later with
    # `make_golden.py --append 3`, which leaves the first three cases and the bootstrap goldens untouched
    ("tiny-llama", shapes.tiny_llama(), 104),
    ("tiny-mistral", shapes.tiny_mistral(), 105),
    # OPT (opt-iml-1.3b) and BLOOM (bloomz-7b1) of the same sweep (:148,151): `make_golden.py --append 5`
    ("tiny-opt", shapes.tiny_opt(), 106),
    ("tiny-bloom", shapes.tiny_bloom(), 107),
    # the Llama block with a q|k|v bias (Qwen of the sweep, :159; HF class Qwen2ForCausalLM): `make_golden.py --append 7`
    ("tiny-qwen", shapes.tiny_qwen(), 108),
]


def scorer_goldens(R, k2_fn, specs=SPECS):
    pert = json.load(open(os.path.join(REF, "data", "perturbations.json")))
    cases = []
    for name, shp, seed in specs:
        W = weights.init_weights(shp, seed=seed)
        model = hf_models.build_hf_model(shp, W)
        tok = HashTokenizer(shp.vocab, eos_token_id=shp.eos_id).apply_pad_default()
        prompts = []
        for q in R.ALL_PROMPTS[:10]:
            prompts.append(("base", f"{FEW_SHOT}Question: {q} Answer either 'Yes' or 'No', without any other text.\nAnswer:"))
        for q in R.ALL_PROMPTS[10:22]:
            prompts.append(("instruct", f"{q} Answer either 'Yes' or 'No', without any other text."))
        for s in range(5):
            for j in (0, 1):
                prompts.append(("perturbation", f"{pert[s]['rephrasings'][j]} {pert[s]['response_format']}"))
        rows = []
        for j, (kind, p) in enumerate(prompts):
            # Every third row pins the candidate pair to a token the model itself generates at
            # position j%7+1, so that look-ahead hits at positions > 0 are exercised (with the
            # hashed " Yes"/" No" ids a random-init model almost never ranks them in the top-k).
            tok.overrides = {}
            cand = None
            if j % 3 == 1:
                ids0 = tok(p, return_tensors="pt").input_ids
                with torch.no_grad():
                    g = model.generate(ids0, max_new_tokens=10, pad_token_id=shp.eos_id, eos_token_id=shp.eos_id)
                gen = g[0][ids0.shape[1]:].tolist()
                a = gen[min(j % 7 + 1, len(gen) - 1)]
                cand = [int(a), int((a + 1) % shp.vocab)]
                if j % 2 == 0:
                    cand = cand[::-1]
                tok.overrides = {" Yes": cand[0], " No": cand[1]}
            with contextlib.redirect_stdout(io.StringIO()):
                r5 = R.get_yes_no_logprobs(model, tok, p, "cpu")
                r2 = k2_fn(model, tok, p, "cpu")
            ids = tok(p).input_ids
            rows.append(dict(kind=kind, prompt=p, n_tokens=len(ids), cand_override=cand,
                             ref_100q={k: r5[k] for k in ("yes_prob", "no_prob", "relative_prob", "completion", "success")},
                             ref_instruct={k: (r2[k] if not isinstance(r2[k], (np.floating,)) else float(r2[k]))
                                           for k in ("yes_logprob", "no_logprob", "yes_prob", "no_prob",
                                                     "relative_prob", "completion", "position_found", "yes_no_found")}))
        cases.append(dict(shape=name, seed=seed, vocab=shp.vocab, eos_id=shp.eos_id, rows=rows))
        print(name, "done:", sum(r["ref_instruct"]["yes_no_found"] for r in rows), "k=2 hits of", len(rows))
    return dict(transformers=__import__("transformers").__version__, torch=torch.__version__, cases=cases)


def bootstrap_goldens(R, scratch):
    out = {}
    sys.path.insert(0, os.path.join(REF, "survey_analysis"))
    with contextlib.redirect_stdout(io.StringIO()):
        import analyze_base_vs_instruct_mae_100q as A

    # (1) synthetic C5-sized case through the reference functions
    rs = np.random.RandomState(1234)
    base = np.abs(rs.normal(0.3, 0.15, 100)).tolist()
    inst = np.abs(rs.normal(0.35, 0.2, 100)).tolist()
    r = A.bootstrap_mae_difference(base, inst)
    out["c5_synthetic"] = dict(base=base, instruct=inst, n_bootstrap=10000, seed=42,
                               expect={k: float(r[k]) for k in ("base_mae", "instruct_mae", "observed_diff",
                                                                 "ci_lower", "ci_upper", "p_value")},
                               first_diffs=r["bootstrap_diffs"][:16].tolist(),
                               sum_diffs=float(np.sum(r["bootstrap_diffs"])))
    fams = {f"F{i}": dict(base=np.abs(rs.normal(0.3, 0.1, 100)).tolist(),
                          instruct=np.abs(rs.normal(0.33, 0.12, 100)).tolist()) for i in range(3)}
    r = A.bootstrap_overall_mae_difference(fams)
    out["c5_overall_synthetic"] = dict(families=fams, expect={k: float(r[k]) for k in (
        "base_mae", "instruct_mae", "observed_diff", "ci_lower", "ci_upper", "p_value")},
        first_diffs=r["bootstrap_diffs"][:16].tolist())

    # (2) shipped data through the reference pipeline; capture the per-family
    #     abs-error vectors the bootstrap consumed so the GPU test can feed the same numbers.
    captured = {}
    orig = A.bootstrap_mae_difference

    def spy(b, i, *a, **k):
        res = orig(b, i, *a, **k)
        captured[len(captured)] = dict(base=list(map(float, b)), instruct=list(map(float, i)),
                                       expect={kk: float(res[kk]) for kk in ("base_mae", "instruct_mae",
                                               "observed_diff", "ci_lower", "ci_upper", "p_value")})
        return res
    A.bootstrap_mae_difference = spy
    orig_over = A.bootstrap_overall_mae_difference
    over = {}

    def spy_over(fe, *a, **k):
        res = orig_over(fe, *a, **k)
        over["families"] = {f: dict(base=list(map(float, e["base"])), instruct=list(map(float, e["instruct"])))
                            for f, e in fe.items()}
        over["expect"] = {kk: float(res[kk]) for kk in ("base_mae", "instruct_mae", "observed_diff",
                                                        "ci_lower", "ci_upper", "p_value")}
        return res
    A.bootstrap_overall_mae_difference = spy_over
    # the script writes results/ next to itself: run on a copy of the module file in scratch
    sa = os.path.join(scratch, "survey_analysis")
    os.makedirs(os.path.join(sa, "results"), exist_ok=True)
    A.BASE_DIR = sa
    if hasattr(A, "RESULTS_DIR"):
        A.RESULTS_DIR = os.path.join(sa, "results")
    with contextlib.redirect_stdout(io.StringIO()):
        try:
            A.main()
        except Exception as e:  # writing outputs may fail; the captured vectors are what we need
            print("main() raised", e, file=sys.stderr)
    out["shipped_families"] = list(captured.values())
    out["shipped_overall"] = over

    # (3) analyze_results(df) of the 100q script on a synthetic 2-family table
    rs = np.random.RandomState(99)
    rows = []
    fam_diffs = {}
    for fam in ("Falcon", "StableLM"):
        b = rs.uniform(0.2, 0.8, 100)
        i = np.clip(b + rs.normal(0.02, 0.1, 100), 0, 1)
        fam_diffs[fam] = (i - b).tolist()
        for q in range(100):
            rows.append(dict(prompt=f"q{q}", model_family=fam, base_or_instruct="base", relative_prob=b[q]))
            rows.append(dict(prompt=f"q{q}", model_family=fam, base_or_instruct="instruct", relative_prob=i[q]))
    df = pd.DataFrame(rows)
    with contextlib.redirect_stdout(io.StringIO()):
        res = R.analyze_results(df)
    out["analyze_results"] = dict(family_diffs=fam_diffs,
                                  expect={f: {k: float(v) for k, v in r.items()} for f, r in res.items()})

    # (4) analyze_results_base_versus_instruct.py (Appendix D.2) with plotting stubbed
    for m in ("matplotlib", "matplotlib.pyplot", "seaborn"):
        sys.modules[m] = MagicMock()
    gd = os.path.join(scratch, "a9", "G:", "My Drive", "Computational", "llm_interpretation")
    os.makedirs(gd, exist_ok=True)
    import shutil
    shutil.copy(os.path.join(REF, "data", "model_comparison_results.csv"), gd)
    os.chdir(os.path.join(scratch, "a9"))
    with contextlib.redirect_stdout(io.StringIO()):
        g = runpy.run_path(os.path.join(REF, "analysis", "analyze_results_base_versus_instruct.py"))
    pv = g["pivot_for_bootstrap"]
    out["a9_shipped"] = dict(pivot=pv.values.tolist(), columns=list(map(str, pv.columns)),
                             expect=dict(mae=float(g["observed_mae"]), mae_ci_lower=float(g["mae_ci_lower"]),
                                         mae_ci_upper=float(g["mae_ci_upper"]), mae_p_value=float(g["p_value_mae"]),
                                         mean_diff=float(g["observed_mean_diff"]),
                                         mean_ci_lower=float(g["mean_diff_ci_lower"]),
                                         mean_ci_upper=float(g["mean_diff_ci_upper"]),
                                         mean_p_value=float(g["p_value_mean_diff_bootstrap"])))
    return out


def main():
    scratch = tempfile.mkdtemp(prefix="golden_")
    R = import_reference_100q(scratch)
    k2 = extract_k2_scorer()
    if len(sys.argv) > 2 and sys.argv[1] == "--append":      # only the cases from SPECS[n:], appended to the file
        n = int(sys.argv[2])
        path = os.path.join(HERE, "scorer_golden.json")
        old = json.load(open(path))
        assert len(old["cases"]) == n, (len(old["cases"]), n)
        old["cases"] += scorer_goldens(R, k2, SPECS[n:])["cases"]
        json.dump(old, open(path, "w"), indent=1)
        print("appended", len(SPECS) - n, "cases")
        return
    sg = scorer_goldens(R, k2)
    json.dump(sg, open(os.path.join(HERE, "scorer_golden.json"), "w"), indent=1)
    bg = bootstrap_goldens(R, scratch)
    json.dump(bg, open(os.path.join(HERE, "bootstrap_golden.json"), "w"), indent=1)
    # the 100 survey questions are an input of the path (hard-coded list, run_base_vs_instruct_100q.py:122-230)
    json.dump(dict(survey1=R.SURVEY1_PROMPTS, survey2=R.SURVEY2_PROMPTS),
              open(os.path.join(HERE, "survey_questions.json"), "w"), indent=1)
    print("wrote goldens")


if __name__ == "__main__":
    main()
