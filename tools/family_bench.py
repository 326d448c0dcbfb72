"""Throughput of the scorer on every decoder-only family of the reference's model lists at FULL size (random-init, FP16 mode,
2,000 perturbation-length prompts, top-5 / 10-position look-ahead; nothing resolves early with random weights, so every prompt
takes all ten positions).  One JSON object per model; the GEMM share comes from the library's own event bracket."""
import argparse
import ctypes as C
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
from llm_interpretation_replication_b200 import _lib, shapes, weights  # noqa: E402
from llm_interpretation_replication_b200.scorer import Scorer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--models", nargs="*", default=["falcon_7b", "redpajama_3b", "stablelm_alpha_7b", "redpajama_7b", "dolly_v2_7b", "mistral_7b",
                                                 "llama2_7b", "qwen_7b", "opt_iml_1_3b", "bloomz_7b1"])
ap.add_argument("--prompts", type=int, default=2000)
ap.add_argument("--mode", default="fp16", choices=["fp16", "bf16_split", "bf16_split_attn32", "bf16"])
args = ap.parse_args()
lib = _lib.load()
MODE = {"fp16": shapes.MODE_FP16, "bf16_split": shapes.MODE_BF16_SPLIT, "bf16_split_attn32": shapes.MODE_BF16_SPLIT_ATTN32,
        "bf16": shapes.MODE_BF16_FAST}[args.mode]
for maker in args.models:
    shp = getattr(shapes, maker)()
    ids, cand, L = bench.synthetic_prompts(shp, args.prompts, seed=2)
    W = weights.init_weights_device(shp, seed=1, device="cuda:0", weight_dtype=torch.float16 if MODE == shapes.MODE_FP16 else torch.bfloat16)
    # batch sized by the scorer from the free device memory (scorer.fit_batch): the KV cache of a multi-head 7B model is 0.5 MB per token
    sc = Scorer(shp, W, mode=MODE, device="cuda:0", max_tokens=None)
    max_tokens = sc.max_tokens
    del W
    sc.score_ids(ids[:300], cand[:300], topk=5, nonfinite="nan")
    torch.cuda.synchronize()
    lib.b2s_launch_count(1)
    lib.b2s_profile_gemm(1)
    t0 = time.time()
    rows, gens = sc.score_ids(ids, cand, topk=5, nonfinite="nan")
    torch.cuda.synchronize()
    dt = time.time() - t0
    ms, n, fl = C.c_double(), C.c_int64(), C.c_double()
    lib.b2s_profile_gemm_read(C.byref(ms), C.byref(n), C.byref(fl))
    lib.b2s_profile_gemm(0)
    params = shp.n_layer * shp.layer_params + shp.vocab * shp.d_model * (1 if shp.tied_head else 2)
    print(json.dumps({"model": shp.name, "layers": shp.n_layer, "params_b": round(params / 1e9, 2), "mode": args.mode, "head_dim": shp.head_dim,
                      "heads": [shp.n_head, shp.n_kv_head], "prompts": len(ids), "prompts_per_s": len(ids) / dt, "seconds": dt,
                      "gemm_share": ms.value / 1e3 / dt, "gemm_tflops": fl.value / (ms.value / 1e3) / 1e12,
                      "launches": int(lib.b2s_launch_count(1)), "batch_tokens": max_tokens, "workspace_gb": round(sc.workspace_bytes / 2**30, 1), "finite": bool(np.isfinite(rows[:, 2]).all())}), flush=True)
    sc.close()
    del sc
    torch.cuda.empty_cache()
