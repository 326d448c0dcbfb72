"""B2S_MODE_INT8_EMU on the GPU: the fake-quantisation kernel bit for bit against its torch restatement, and the scorer in that
mode against the oracle that restates bitsandbytes' Linear8bitLt call by call (oracle/int8_ref.py; parity against the
bitsandbytes library itself is unpinned: it is not in this image)."""
import json
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _fakequant_torch(x, q_start, q_len, thr):
    out = x.float().clone()
    for s, n in zip(q_start, q_len):
        a = x[s:s + n].float()
        big = a.abs() >= thr
        idx = big.any(0)
        sca = torch.where(big, torch.zeros_like(a), a.abs()).amax(1, keepdim=True)
        q = torch.round(a * (torch.full_like(sca, 127.0) / sca)) * (sca / 127.0)      # (tensor / tensor: correctly rounded, as the kernel's division)
        q = torch.where(sca > 0, q, a)
        q[:, idx] = a[:, idx]
        out[s:s + n] = q
    return out.half()


@pytest.mark.parametrize("K", [192, 4544])
def test_int8_fakequant_kernel_bit_exact(K):
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(K)
    lens = [1, 37, 128, 5, 1, 64]
    T = sum(lens)
    x = (torch.randn(T, K, generator=g) * 1.7).half()
    x[3, 10] = 7.0; x[40, K - 1] = -6.0; x[41, 0] = 6.5; x[170, 77] = 30.0       # outliers in prompts 1, 2 and 3 (row 170)
    x[200] = 0                                                                     # an all-zero row
    q_start = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int32)
    q_len = np.array(lens, np.int32)
    want = _fakequant_torch(x, q_start.tolist(), lens, 6.0)
    xd = x.to(DEV).contiguous()
    mask = torch.empty(len(lens) * K, dtype=torch.uint8, device=DEV)
    d_start, d_len = torch.from_numpy(q_start).to(DEV), torch.from_numpy(q_len).to(DEV)      # (named: they must outlive the call)
    _lib.check(lib.b2s_int8_fakequant_f16(_lib.ptr(xd), T, K, _lib.ptr(d_start), _lib.ptr(d_len), len(lens), _lib.ptr(mask), 6.0,
                                          _lib.current_stream()), "int8_fakequant")
    torch.cuda.synchronize()
    got = xd.cpu()
    bad = (got != want).nonzero()
    assert bad.numel() == 0, (bad.shape[0], [(int(r), int(c), float(x[r, c]), float(got[r, c]), float(want[r, c])) for r, c in bad[:6]])
    assert not torch.equal(got, x)                           # something was quantised
    assert torch.equal(got[:, 10][1:38], x[:, 10][1:38])     # the outlier column of prompt 1 stayed fp16 for all of its rows
    m = mask.cpu().view(len(lens), K).bool()
    want_mask = torch.stack([(x[s:s + n].float().abs() >= 6.0).any(0) for s, n in zip(q_start.tolist(), lens)])
    assert torch.equal(m, want_mask) and m[1, 10] and m[3, 77] and m[2, 0] and m[2, K - 1]
    # threshold 0 = no decomposition: every element goes on the grid of its row's absmax
    xd0 = x.to(DEV).contiguous()
    _lib.check(lib.b2s_int8_fakequant_f16(_lib.ptr(xd0), T, K, _lib.ptr(d_start), _lib.ptr(d_len), len(lens), _lib.ptr(mask), 0.0,
                                          _lib.current_stream()), "int8_fakequant")
    torch.cuda.synchronize()
    assert torch.equal(xd0.cpu(), _fakequant_torch(x, q_start.tolist(), lens, float("inf")))


def _with_outliers(W, shp):
    """Push a few LayerNorm gains up so that |x| >= 6 occurs in the inputs of the linear layers (random-init models have none)."""
    W = dict(W)
    for l in range(shp.n_layer):
        g = W[f"l{l}.ln1_g"].clone()
        g[5 + l] = 9.0
        g[17] = -8.0
        W[f"l{l}.ln1_g"] = g
    return W


@pytest.mark.parametrize("name", ["tiny-falcon", "tiny-neox-seq", "tiny-llama"])
def test_int8_emulation_vs_restated_linear8bitlt(name):
    """Prompts of the golden file, look-ahead included, against decoder_ref with every block's linear layer replaced by the
    restated Linear8bitLt (one call per prompt pass, one per generated token).  The two differ by fp16 operand rounding only;
    the int8 grid itself moves relative_prob by far more, which the last assertion shows."""
    import gpu_util as U
    from llm_interpretation_replication_b200 import shapes, weights
    from llm_interpretation_replication_b200 import scorer as S
    from oracle import decoder_ref, int8_ref
    case = [c for c in U.golden_cases() if c["shape"] == name][0]
    shp = shapes.REGISTRY[name]()
    W = _with_outliers(weights.init_weights(shp, seed=case["seed"]), shp)
    f16 = lambda t: t.to(torch.float16).to(t.dtype)
    rows_in = case["rows"][:14]
    ids, cands = [], []
    for row in rows_in:
        tok = U.tok_for(shp, row)
        ids.append(tok(row["prompt"]).input_ids)
        cands.append(S.candidate_ids(tok, "100q"))
    cands = np.array(cands)
    sc8 = U.make_scorer(shp, W, shapes.MODE_INT8_EMU, max_tokens=4096, max_seqs=64)
    r8, g8 = sc8.score_ids(ids, cands, topk=5, eos_id=shp.eos_id)
    sc8.close()
    sc16 = U.make_scorer(shp, W, shapes.MODE_FP16, max_tokens=4096, max_seqs=64)
    r16, _ = sc16.score_ids(ids, cands, topk=5, eos_id=shp.eos_id)
    sc16.close()
    errs, shifts, flips = [], [], 0
    for i, x in enumerate(ids):
        o8 = decoder_ref.score_ids(shp, W, x, int(cands[i][0]), int(cands[i][1]), topk=5, eos_id=shp.eos_id, round_act=f16,
                                   linear_fn=int8_ref.make_linear_fn(f16, prompt_len=len(x)))
        o16 = decoder_ref.score_ids(shp, W, x, int(cands[i][0]), int(cands[i][1]), topk=5, eos_id=shp.eos_id, round_act=f16)
        if o8["pos_found"] != int(r8[i][6]):
            flips += 1
            continue
        errs.append(abs(float(r8[i][2]) - o8["rel"]))
        if o16["pos_found"] == o8["pos_found"]:
            shifts.append(abs(o8["rel"] - o16["rel"]))
    rec = {"case": name, "rows": len(ids), "position_mismatches": flips,
           "emulation_vs_restated_int8_abs_delta_rel_prob": {"mean": float(np.mean(errs)), "max": float(np.max(errs))},
           "restated_int8_vs_fp16_abs_delta_rel_prob": {"mean": float(np.mean(shifts)), "max": float(np.max(shifts))}}
    if os.path.isdir("gpurun_out"):
        with open("gpurun_out/int8_parity.jsonl", "a") as f:
            f.write(json.dumps(rec) + "\n")
    # rint() turns a one-ulp difference of an fp16 input into a whole int8 step for every element that sits near a rounding
    # boundary (about ulp / step = 3 % of them), so LLM.int8 agrees with ITSELF only statistically once the inputs come from
    # another correct implementation: tests/test_int8_cpu.py measures that self-disagreement (1.1e-3 mean on these prompts, as
    # large as the int8-vs-fp16 shift).  The emulation has to sit at that level, not below it; the kernel itself is bit-exact
    # on identical inputs (test above).
    assert flips <= 2, rec
    assert np.max(errs) <= 1e-2 and np.mean(errs) <= 1.6 * np.mean(shifts) + 2e-4, rec
    assert not np.array_equal(r8[:, 2], r16[:, 2])                                  # the mode is not a no-op


def test_int8_emulation_real_width():
    """Falcon-7B widths at 2 layers (d 4544, 71 heads, MQA): the emulation against the restated Linear8bitLt."""
    import gpu_util as U
    from llm_interpretation_replication_b200 import shapes, weights
    from oracle import decoder_ref, int8_ref
    shp = shapes.falcon_7b().with_layers(2)
    W = _with_outliers(weights.init_weights(shp, seed=23), shp)
    f16 = lambda t: t.to(torch.float16).to(t.dtype)
    g = np.random.RandomState(8)
    ids = [g.randint(3, shp.vocab - 3, size=L).tolist() for L in (29, 118, 66)]
    sc = U.make_scorer(shp, W, shapes.MODE_INT8_EMU, max_tokens=1024, max_seqs=8, max_lookahead=2)
    rows, gens = sc.score_ids(ids, np.array([[700, 900]]), topk=5, max_lookahead=2)
    sc.close()
    worst = 0.0
    for i, x in enumerate(ids):
        lo = decoder_ref.last_logits(shp, W, torch.tensor(x), round_act=f16, linear_fn=int8_ref.make_linear_fn(f16, prompt_len=len(x)))
        p = torch.softmax(lo, -1)
        worst = max(worst, abs(float(rows[i][2]) - (p[700] / (p[700] + p[900])).item()))
    if os.path.isdir("gpurun_out"):
        with open("gpurun_out/int8_parity.jsonl", "a") as f:
            f.write(json.dumps({"case": "falcon-7b widths, 2 layers", "rows": 3, "worst_abs_delta_rel_prob_vs_restated_int8": worst}) + "\n")
    assert worst <= 2e-2, worst      # (see the statistical bound above: boundary elements flip by whole int8 steps)


def test_int8_mode_is_refused_for_encoder_decoder_models():
    from llm_interpretation_replication_b200 import shapes, t5, _lib
    shp = t5.tiny_t5()
    with pytest.raises((_lib.B2SError, AssertionError)):
        t5.EncDecScorer(shp, t5.to_device(t5.init_weights(shp, 1), DEV, torch.float16), mode=shapes.MODE_INT8_EMU, device=DEV)
