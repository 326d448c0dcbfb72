"""Full-depth parity: the BASELINE configs themselves, at their real depth, against the fp32 HF oracle.

  C3  Falcon-7B, 32 layers, perturbation-length prompts (>= 4 from each of [64,96) / [96,128) / [128,160) / [160,192)),
      top-5 / 10-position look-ahead, FP16 (the default 16-bit mode), BF16_FAST and BF16_SPLIT;
  C2  StableLM-base-alpha-7B, 16 layers, survey questions in the base (few-shot) and instruct formats, FP16 / BF16_FAST;
  C1  RedPajama-INCITE-Base-3B, 32 layers, survey questions in the base format, FP32_CHECK at <= 1e-5 (and FP16).

Oracle = oracle/scorer_oracle.get_yes_no_logprobs (the reference's function restated, pinned to goldens produced by
the reference itself) over HF transformers fp32 on the host cores, with the SAME bf16-representable weights widened
to fp32 and max_new_tokens = the 10 scanned positions (positions 10..49 of the reference call only feed the completion
string).  Gates (BASELINE.json north_star): |delta relative_prob| <= 2e-3 per prompt in the 16-bit mode, <= 1e-5 in the
fp32 check mode, identical look-ahead position.  The position is a discontinuous function of the logits (a greedy
token or a top-k membership can flip on a 1e-3 logit difference), so in the 16-bit modes a disagreement is COUNTED and
bounded (at most one prompt per config here), never skipped silently; the fp32 mode allows none.  BF16_FAST sits at the bf16 noise floor at this depth:
it is measured and bounded here, not held to 2e-3 per prompt; DESIGN.md states the pass rates.
Every run appends its numbers to gpurun_out/full_depth_parity.jsonl when that directory exists.

Sizes are bounded so the whole file takes a few minutes of host time (HF fp32 on the box's cores is the slow part);
B2S_FULL_DEPTH_SCALE=k multiplies the number of prompts.
"""
import gc
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

SCALE = int(os.environ.get("B2S_FULL_DEPTH_SCALE", "1"))
TOL_16 = 2e-3
TOL_FP32 = 1e-5
LOOKAHEAD = 10
# BF16_FAST is NOT held to the gate: measured max 7.1e-3 / 80 % of 1,000 Falcon-7B prompts within 2e-3
# (profiles/r2_parity_histogram.json) and 7.5e-3 on StableLM-7B.  The bound only catches a regression of the mode.
BF16_FAST_BOUND = 1.5e-2


def _letters_prompt(rs, n_tokens):
    letters = np.array(list("abcdefghijklmnopqrstuvwxyz ABCDEFGHIJKLMNOPQRSTUVWXYZ,.?'"))
    return "".join(rs.choice(letters, size=4 * n_tokens))          # 4 bytes per token under the hash tokenizer


def _log(rec):
    if os.path.isdir("gpurun_out"):
        with open("gpurun_out/full_depth_parity.jsonl", "a") as f:
            f.write(json.dumps(rec) + "\n")
    print(json.dumps(rec))


def _run_config(name, shp, prompts, topk, variant, modes, seed, n_hits):
    """prompts: strings.  modes: {label: (mode, tol or None, max position mismatches)}.  Returns per-mode stats."""
    import gpu_util as U
    from llm_interpretation_replication_b200 import shapes, weights
    from llm_interpretation_replication_b200.scorer import Scorer
    from llm_interpretation_replication_b200.tokenizer import HashTokenizer
    from oracle import hf_models, scorer_oracle
    import transformers
    transformers.logging.set_verbosity_error()
    dev = torch.device(U.DEV)
    W = weights.init_weights_device(shp, seed=seed, device=dev, weight_dtype=torch.bfloat16, randomize_affine=True)
    tok = HashTokenizer(shp.vocab, eos_token_id=shp.eos_id).apply_pad_default()
    ids = [tok(p).input_ids for p in prompts]
    rs = np.random.RandomState(seed)
    cands = []
    for _ in prompts:
        a, b = rs.choice(np.arange(64, shp.vocab - 64), size=2, replace=False)
        cands.append([int(a), int(b)])

    # ---- oracle: fp32 HF on the host, same weights ------------------------------------------------------------
    torch.set_num_threads(os.cpu_count() or 1)
    Wc = {k: v.detach().to("cpu", torch.float32) for k, v in W.items() if k not in ("rope_cos", "rope_sin")}
    if shp.tied_head:
        Wc["lm_head"] = Wc["wte"]
    model = hf_models.build_hf_model(shp, Wc)
    want = []
    for k, p in enumerate(prompts):
        o = scorer_oracle.get_yes_no_logprobs(model, tok, p, variant=variant, cand_ids=tuple(cands[k]), max_new_tokens=LOOKAHEAD)
        if k < n_hits:
            # make this prompt resolve at a later position: candidate a = the token greedy decoding emits there
            pos = 1 + (k % 4)
            gen = o["_gen_ids"]
            if len(gen) > pos and gen[pos] != cands[k][1]:
                cands[k][0] = int(gen[pos])
                o = scorer_oracle.get_yes_no_logprobs(model, tok, p, variant=variant, cand_ids=tuple(cands[k]),
                                                      max_new_tokens=LOOKAHEAD)
        want.append(o)
    del model, Wc
    gc.collect()
    pos_key = "_position_found" if variant == "100q" else "position_found"
    found_key = "_found" if variant == "100q" else "yes_no_found"
    assert sum(1 for o in want if o[found_key]) >= min(n_hits, 1), "no prompt resolves at a later position: the case is too easy"

    # ---- CUDA path, every mode ------------------------------------------------------------------------------------
    stats = {}
    total = sum(len(x) for x in ids)
    for label, (mode, tol, max_pos_mismatch) in modes.items():
        Wm = W if mode != shapes.MODE_FP32_CHECK else {k: (v.float() if weights.is_matrix(k) else v) for k, v in W.items()}
        if mode == shapes.MODE_FP32_CHECK and shp.tied_head:
            Wm["lm_head"] = Wm["wte"]
        sc = Scorer(shp, Wm, mode=mode, device=dev, max_tokens=max(total, 256), max_seqs=len(ids), max_lookahead=LOOKAHEAD)
        rows, gens = sc.score_ids(ids, np.array(cands), topk=topk, max_lookahead=LOOKAHEAD, eos_id=shp.eos_id)
        sc.close()
        del sc, Wm
        torch.cuda.empty_cache()
        errs, mismatches = [], []
        for k, o in enumerate(want):
            if int(rows[k][6]) != o[pos_key] or bool(rows[k][7] > 0) != bool(o[found_key]):
                mismatches.append(k)                     # counted and bounded, never skipped silently
                continue
            ref_rel = o["relative_prob"] if variant != "base_vs" else o["yes_prob"] / (o["yes_prob"] + o["no_prob"])
            errs.append(abs(float(rows[k][2]) - ref_rel))
            assert abs(float(rows[k][0]) - o["yes_prob"]) <= max(tol or 1.0, 1e-5) and abs(float(rows[k][1]) - o["no_prob"]) <= max(tol or 1.0, 1e-5)
            if mode == shapes.MODE_FP32_CHECK:           # identical greedy tokens over the positions this prompt needed
                got = [int(t) for t in gens[k] if t >= 0]
                assert got == o["_gen_ids"][:len(got)], (label, k)
        e = np.array(errs) if errs else np.zeros(1)
        rec = {"config": name, "mode": label, "n_layer": shp.n_layer, "prompts": len(prompts), "tokens": total,
               "max_abs_delta_rel_prob": float(e.max()), "mean_abs_delta_rel_prob": float(e.mean()),
               "pass_rate_2e-3": float((e <= TOL_16).mean()), "position_mismatches": len(mismatches),
               "resolved_later": int(sum(1 for o in want if o[found_key] and o[pos_key] > 0))}
        _log(rec)
        stats[label] = rec
        assert len(mismatches) <= max_pos_mismatch, (label, mismatches)
        if tol is not None:
            assert e.max() <= tol, (label, rec)
    return stats


def test_c3_falcon_7b_32_layers_length_stratified():
    """BASELINE configs[2]: Falcon-7B x 32 layers; 4 prompts from each perturbation length range; 4 of the 16 resolve at a
    later look-ahead position."""
    from llm_interpretation_replication_b200 import shapes
    shp = shapes.falcon_7b()
    rs = np.random.RandomState(3)
    lens = []
    for lo, hi in ((64, 96), (96, 128), (128, 160), (160, 192)):
        lens += [int(x) for x in rs.randint(lo, hi, size=4 * SCALE)]
    order = np.argsort(rs.rand(len(lens)))                     # hits spread over the length ranges
    prompts = [_letters_prompt(rs, lens[i]) for i in order]
    st = _run_config("C3 falcon-7b", shp, prompts, topk=5, variant="100q", seed=0, n_hits=4 * SCALE,
                     modes={"fp16": (shapes.MODE_FP16, TOL_16, 1),
                            # bf16 operands: noise floor at this depth -- bounded, pass rate logged
                            "bf16_fast": (shapes.MODE_BF16_FAST, BF16_FAST_BOUND, 1),
                            # bf16 weights untouched, hi+lo activations: holds the gate per prompt
                            "bf16_split": (shapes.MODE_BF16_SPLIT, TOL_16, 1)})
    assert st["fp16"]["pass_rate_2e-3"] == 1.0 and st["bf16_split"]["pass_rate_2e-3"] == 1.0


def test_c2_stablelm_alpha_7b_16_layers_both_formats():
    """BASELINE configs[1]: StableLM-base-alpha-7B (GPT-NeoX, parallel residual, partial rotary, biases) x 16 layers on the
    survey questions in the base (few-shot, ~80 tokens) and the instruct (bare, ~25 tokens) formats."""
    from llm_interpretation_replication_b200 import shapes
    from llm_interpretation_replication_b200.entrypoints import common
    shp = shapes.stablelm_alpha_7b()
    s1, s2 = common.survey_questions()
    qs = (s1 + s2)[::17][:6 * SCALE]
    prompts = [common.format_base(q) for q in qs] + [common.format_instruct(q) for q in qs]
    _run_config("C2 stablelm-alpha-7b", shp, prompts, topk=5, variant="100q", seed=1, n_hits=3 * SCALE,
                modes={"fp16": (shapes.MODE_FP16, TOL_16, 1), "bf16_fast": (shapes.MODE_BF16_FAST, BF16_FAST_BOUND, 1),
                       "bf16_split": (shapes.MODE_BF16_SPLIT, TOL_16, 1)})


def test_c1_redpajama_3b_32_layers_fp32_check():
    """BASELINE configs[0]: RedPajama-INCITE-Base-3B (GPT-NeoX, head_dim 80, sequential residual) x 32 layers in the fp32
    check mode, gate 1e-5, on survey questions in the base format -- the reference's own CPU-runnable case."""
    from llm_interpretation_replication_b200 import shapes
    from llm_interpretation_replication_b200.entrypoints import common
    shp = shapes.redpajama_3b()
    s1, s2 = common.survey_questions()
    qs = (s1 + s2)[5::13][:6 * SCALE]
    prompts = [common.format_base(q) for q in qs]
    _run_config("C1 redpajama-3b", shp, prompts, topk=5, variant="100q", seed=2, n_hits=2 * SCALE,
                modes={"fp32_check": (shapes.MODE_FP32_CHECK, TOL_FP32, 0), "fp16": (shapes.MODE_FP16, TOL_16, 1)})
