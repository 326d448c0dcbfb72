"""Transitive full-depth parity over thousands of prompts, all on the GPU.

The fp32 check mode is pinned to the fp32 HF oracle at <= 1e-5 (tests/test_gpu_scorer.py goldens, and
tests/test_gpu_full_depth.py at the real depth of the BASELINE configs); here it is the yardstick for the 16-bit
modes on a large sample of the bench prompt set: Falcon-7B, 32 layers, top-5 / 10-position look-ahead, same
bf16-representable weights in every mode.  Reports, per mode, the |delta relative_prob| histogram, the pass rate at the
2e-3 gate, and how often the look-ahead position or the found flag disagrees (those prompts are counted, not compared).

  python tools/parity_histogram.py [n_prompts=2000] [n_layer=32] [modes=fp16,bf16,bf16_split] [shape=falcon_7b] > gpurun_out/parity_histogram.json
(shape: any maker of shapes.py, e.g. llama2_7b -- multi-head, head_dim 128: the tensor-core prompt pass at 128, the streaming
one-token attention and the vectorised RoPE kernel; n_layer 0 = the shape's own depth)
"""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import shapes, weights  # noqa: E402
from llm_interpretation_replication_b200.scorer import Scorer  # noqa: E402
import bench  # noqa: E402

EDGES = [0, 1e-6, 3e-6, 1e-5, 3e-5, 1e-4, 3e-4, 1e-3, 2e-3, 3e-3, 5e-3, 1e-2, 1.0]


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
    n_layer = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    which = sys.argv[3].split(",") if len(sys.argv) > 3 else ["fp16", "bf16", "bf16_split"]
    modes = {"fp16": shapes.MODE_FP16, "bf16": shapes.MODE_BF16_FAST, "bf16_split": shapes.MODE_BF16_SPLIT,
             "bf16_split_attn32": shapes.MODE_BF16_SPLIT_ATTN32}
    shp = getattr(shapes, sys.argv[4])() if len(sys.argv) > 4 else shapes.falcon_7b()
    if n_layer and n_layer != shp.n_layer:
        shp = shp.with_layers(n_layer)
    dev = torch.device("cuda:0")
    import os
    from llm_interpretation_replication_b200 import _lib
    for kv in filter(None, os.environ.get("B2S_OPTIONS", "").split(",")):      # e.g. B2S_OPTIONS=attn_ts=0 : A/B of one kernel choice
        k, v = kv.split("=")
        _lib.check(_lib.load().b2s_set_option(k.encode(), int(v)), "set_option")
    W = weights.init_weights_device(shp, seed=0, device=dev, weight_dtype=torch.bfloat16, randomize_affine=True)
    ids, cand, L = bench.synthetic_prompts(shp, 10000)
    pick = np.linspace(0, 9999, n).astype(int)
    my_ids = [ids[i].tolist() for i in pick]
    my_cand = cand[pick].copy()
    out = {"shape": shp.name, "n_layer": shp.n_layer, "prompts": n, "tokens": int(sum(len(x) for x in my_ids)),
           "topk": 5, "max_lookahead": 10, "yardstick": "FP32_CHECK mode of the same library (fp32 weights = the bf16 weights widened)",
           "edges": EDGES, "modes": {}}

    def run(mode, Wm, cands, max_tokens):
        sc = Scorer(shp, Wm, mode=mode, device=dev, max_tokens=max_tokens, max_seqs=1024, max_lookahead=10)
        t0 = time.time()
        rows, gens = sc.score_ids(my_ids, cands, topk=5, max_lookahead=10)
        dt = time.time() - t0
        sc.close()
        del sc
        torch.cuda.empty_cache()
        return rows, gens, dt

    W32 = {k: (v.float() if weights.is_matrix(k) else v) for k, v in W.items() if k != "lm_head"}
    W32["lm_head"] = W32["wte"] if shp.tied_head else W["lm_head"].float()
    # pass 1 (fp32): the greedy tokens; a tenth of the prompts then gets candidate a = the token emitted at position
    # 1 + (i mod 8), so that the look-ahead rule is exercised at later positions too, as in the goldens
    _, g32, _ = run(shapes.MODE_FP32_CHECK, W32, my_cand, 32768)
    for i in range(0, n, 10):
        p = 1 + (i // 10) % 8
        if g32[i][p] >= 0 and g32[i][p] != my_cand[i][1]:
            my_cand[i][0] = g32[i][p]
    ref, gref, dt32 = run(shapes.MODE_FP32_CHECK, W32, my_cand, 32768)
    del W32
    torch.cuda.empty_cache()
    out["fp32_check_seconds"] = dt32
    out["resolved_at_later_position_fp32"] = int((ref[:, 6] > 0).sum())
    for name in which:
        rows, gens, dt = run(modes[name], W, my_cand, 65536)
        same = (rows[:, 6] == ref[:, 6]) & ((rows[:, 7] > 0) == (ref[:, 7] > 0))
        e = np.abs(rows[same, 2].astype(np.float64) - ref[same, 2].astype(np.float64))
        hist, _ = np.histogram(e, bins=EDGES)
        first_tok_differs = int((gens[:, 0] != gref[:, 0]).sum())
        out["modes"][name] = {
            "seconds": dt, "compared": int(same.sum()), "position_or_found_disagreements": int((~same).sum()),
            "greedy_token_0_differs": first_tok_differs, "nonfinite_rows": int((~np.isfinite(rows[:, 2])).sum()),
            "max": float(e.max()), "mean": float(e.mean()), "p50": float(np.percentile(e, 50)), "p99": float(np.percentile(e, 99)),
            "p999": float(np.percentile(e, 99.9)), "pass_rate_2e-3": float((e <= 2e-3).mean()), "histogram": hist.tolist()}
        print(json.dumps({name: out["modes"][name]}), file=sys.stderr, flush=True)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
