"""Throughput of the encoder-decoder scorer at the size of the reference's two encoder-decoder sweep models
(bigscience/T0_3B, allenai/tk-instruct-3b-def: t5-xl-lm-adapt, 24 + 24 layers, d_model 2048; reference
compare_instruct_models.py:147,150), random-init, prompts with the perturbation length histogram (</s> appended),
top-2 / 10-position look-ahead as that script scans.  One JSON object per mode; `--shape t5-v1_1-base` for the pair of
compare_base_vs_instruct.py:137-138.  The GEMM share comes from the library's own event bracket (b2s_profile_gemm)."""
import argparse
import ctypes as C
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
from llm_interpretation_replication_b200 import _lib, shapes, t5  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="t5-xl-lm-adapt")
    ap.add_argument("--prompts", type=int, default=2000)
    ap.add_argument("--modes", nargs="*", default=["bf16_split", "fp16", "bf16"])
    ap.add_argument("--max-tokens", type=int, default=32768)
    ap.add_argument("--max-seqs", type=int, default=1024)
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--steps", type=int, default=2)
    args = ap.parse_args()
    shp = t5.REGISTRY[args.shape]()
    if args.layers:
        from dataclasses import replace
        shp = replace(shp, n_enc_layer=args.layers, n_dec_layer=args.layers)
    L = bench.synthetic_lengths(args.prompts, seed=3)
    rs = np.random.RandomState(4)
    ids = [np.concatenate([rs.randint(32, shp.vocab - 32, size=int(l) - 1), [shp.eos_id]]).astype(np.int32) for l in L]
    cand = np.tile(np.array([[1001, 1002]], np.int32), (len(ids), 1))
    lib = _lib.load()
    modes = {"bf16_split": shapes.MODE_BF16_SPLIT, "fp16": shapes.MODE_FP16, "bf16": shapes.MODE_BF16_FAST,
             "fp16_ffn_split": shapes.MODE_FP16_FFN_SPLIT}
    for name in args.modes:
        mode = modes[name]
        dt = torch.bfloat16      # (the scorer converts to what the mode needs; FP16_FFN_SPLIT keeps wo in bf16)
        W = t5.init_weights(shp, seed=1, device="cuda:0", weight_dtype=dt)
        sc = t5.EncDecScorer(shp, W, mode=mode, device="cuda:0", max_tokens=args.max_tokens, max_seqs=args.max_seqs)
        del W
        rows, gens = sc.score_ids(ids[:256], cand[:256], topk=2, eos_id=shp.eos_id, nonfinite="nan")     # warm-up
        torch.cuda.synchronize()
        lib.b2s_launch_count(1)
        lib.b2s_profile_gemm(1)
        t0 = time.time()
        for _ in range(args.steps):
            rows, gens = sc.score_ids(ids, cand, topk=2, eos_id=shp.eos_id, nonfinite="nan")
        torch.cuda.synchronize()
        dt_s = (time.time() - t0) / args.steps
        ms, n, fl = C.c_double(), C.c_int64(), C.c_double()
        lib.b2s_profile_gemm_read(C.byref(ms), C.byref(n), C.byref(fl))
        lib.b2s_profile_gemm(0)
        launches = lib.b2s_launch_count(1) // args.steps
        steps_taken = (gens >= 0).sum(axis=1)
        print(json.dumps({"shape": shp.name, "layers": [shp.n_enc_layer, shp.n_dec_layer], "mode": name, "prompts": len(ids),
                          "prompt_tokens": int(L.sum()), "seconds_per_pass": dt_s, "prompts_per_s": len(ids) / dt_s,
                          "gemm_ms_per_pass": ms.value / args.steps, "gemm_share": ms.value / args.steps / 1e3 / dt_s,
                          "gemm_tflops": fl.value / (ms.value / 1e3) / 1e12 if ms.value > 0 else None,
                          "launches_per_pass": int(launches), "mean_decoder_positions": float(steps_taken.mean()),
                          "finite": bool(np.isfinite(rows[:, 2]).all()), "workspace_gb": sc.workspace_bytes / 2**30, "max_tokens": args.max_tokens}), flush=True)
        sc.close()
        del sc
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
