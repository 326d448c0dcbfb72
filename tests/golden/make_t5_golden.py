#!/usr/bin/env python
"""Generate tests/golden/t5_golden.json by running the REFERENCE's own encoder-decoder branch.

Run in the build container only (needs /root/reference, read-only):
    python tests/golden/make_t5_golden.py
What is executed from the reference (the `is_encoder_decoder` branch of each get_yes_no_logprobs):
  * analysis/run_base_vs_instruct_100q.py:290-325        (top-5 scan, max_new_tokens=50)      -> ref_100q
  * analysis/compare_instruct_models.py:178-226          (top-2 scan, max_new_tokens=50)      -> ref_instruct
  * analysis/compare_base_vs_instruct.py:192-237         (top-2 scan, max_length=50)          -> ref_base_vs
driven with random-init HF ``T5ForConditionalGeneration`` models built from llm_interpretation_replication_b200.t5
(seeded, so the tests rebuild the identical weights) and the offline HashTokenizer in its T5 form (every encoded string
closed with </s> = 1, pad = 0).  Nothing here is read at test time except the JSON output.
"""
import ast
import contextlib
import io
import json
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
REF = "/root/reference"

import make_golden as MG  # noqa: E402  (import_reference_100q, FEW_SHOT)
from llm_interpretation_replication_b200 import t5  # noqa: E402
from llm_interpretation_replication_b200.tokenizer import HashTokenizer  # noqa: E402
from oracle import hf_models  # noqa: E402


def extract_scorer(script):
    src = open(os.path.join(REF, "analysis", script)).read()
    tree = ast.parse(src)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "get_yes_no_logprobs"][0]
    ns = {"torch": torch, "np": np, "MAX_LOOK_AHEAD": 10}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), script, "exec"), ns)
    return ns["get_yes_no_logprobs"]


SPECS = [("tiny-t5", t5.tiny_t5(), 201), ("tiny-t5-relu", t5.tiny_t5_relu(), 202)]


def t5_tokenizer(shp):
    return HashTokenizer(shp.vocab, eos_token_id=shp.eos_id, append_eos=True, pad_token_id=shp.pad_id)


def main():
    scratch = tempfile.mkdtemp()
    R = MG.import_reference_100q(scratch)
    k2_fn = extract_scorer("compare_instruct_models.py")
    bv_fn = extract_scorer("compare_base_vs_instruct.py")
    pert = json.load(open(os.path.join(REF, "data", "perturbations.json")))
    cases = []
    for name, shp, seed in SPECS:
        W = t5.init_weights(shp, seed=seed)
        model = hf_models.build_hf_t5(shp, W)
        tok = t5_tokenizer(shp)
        prompts = []
        for q in R.ALL_PROMPTS[:8]:
            prompts.append(("base", f"{MG.FEW_SHOT}Question: {q} Answer either 'Yes' or 'No', without any other text.\nAnswer:"))
        for q in R.ALL_PROMPTS[10:20]:
            prompts.append(("instruct", f"{q} Answer either 'Yes' or 'No', without any other text."))
        for s in range(1, 5):
            for j in (0, 1):
                prompts.append(("perturbation", f"{pert[s]['rephrasings'][j]} {pert[s]['response_format']}"))
        rows = []
        for j, (kind, p) in enumerate(prompts):
            # Two rows of three pin the candidate pair to a token the model itself generates at position j % 7, so that
            # look-ahead hits at positions > 0 occur (hashed "Yes"/"No" ids almost never rank in a random model's top-k)
            tok.overrides = {}
            cand = None
            if j % 3 != 0:
                ids0 = tok(p, return_tensors="pt").input_ids
                with torch.no_grad():
                    g = model.generate(ids0, max_new_tokens=10, pad_token_id=shp.pad_id, eos_token_id=shp.eos_id, do_sample=False)
                gen = g[0][1:].tolist()
                a = gen[min(j % 7, len(gen) - 1)]
                cand = [int(a), int((a + 1) % shp.vocab)]
                if j % 2 == 0:
                    cand = cand[::-1]
                tok.overrides = {"Yes": cand[0], "No": cand[1]}
            with contextlib.redirect_stdout(io.StringIO()):
                r5 = R.get_yes_no_logprobs(model, tok, p, "cpu")
                r2 = k2_fn(model, tok, p, "cpu")
                rb = bv_fn(model, tok, p, "cpu")
            f = lambda v: float(v) if isinstance(v, (np.floating, float)) else v
            rows.append(dict(kind=kind, prompt=p, n_tokens=len(tok(p).input_ids), cand_override=cand,
                             ref_100q={k: f(r5[k]) for k in ("yes_prob", "no_prob", "relative_prob", "completion", "success")},
                             ref_instruct={k: f(r2[k]) for k in ("yes_logprob", "no_logprob", "yes_prob", "no_prob", "relative_prob",
                                                                  "completion", "position_found", "yes_no_found")},
                             ref_base_vs={k: f(rb[k]) for k in ("yes_logprob", "no_logprob", "yes_prob", "no_prob", "odds_ratio",
                                                                 "completion", "position_found", "yes_no_found")}))
        cases.append(dict(shape=name, seed=seed, vocab=shp.vocab, eos_id=shp.eos_id, rows=rows))
        print(name, "done:", sum(r["ref_instruct"]["yes_no_found"] for r in rows), "k=2 hits of", len(rows),
              "positions", sorted(set(r["ref_instruct"]["position_found"] for r in rows)))
    out = dict(transformers=__import__("transformers").__version__, torch=torch.__version__, cases=cases)
    json.dump(out, open(os.path.join(HERE, "t5_golden.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
