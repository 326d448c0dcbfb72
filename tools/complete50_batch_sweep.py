"""The 50-token completion call (scan 10 positions, generate 50 tokens) on the bench workload at three batch sizes: how much the
one-token steps gain from more rows per step (profiles/r2_bench_batch_tokens.jsonl: 511 / 521 / 533 prompts/s at 5 / 2 / 1 batches)."""
import sys, time, json, numpy as np, torch
sys.path.insert(0, ".")
import bench
from llm_interpretation_replication_b200 import shapes, weights
from llm_interpretation_replication_b200.scorer import Scorer
shp = shapes.falcon_7b()
ids, cand, L = bench.synthetic_prompts(shp, 10000, seed=0)
W = weights.init_weights_device(shp, seed=0, device="cuda:0", weight_dtype=torch.float16)
for bt in (262144, 600000, 1200000):
    sc = Scorer(shp, W, mode=shapes.MODE_FP16, device="cuda:0", max_tokens=bt, max_seqs=max(2048, bt // 64), max_lookahead=50)
    sc.score_ids(ids[:500], cand[:500], topk=5, gen_tokens=50, nonfinite="nan")
    torch.cuda.synchronize(); t0 = time.time()
    rows, gens = sc.score_ids(ids, cand, topk=5, gen_tokens=50, nonfinite="nan")
    torch.cuda.synchronize(); dt = time.time() - t0
    print(json.dumps({"batch_tokens": bt, "complete50_prompts_per_s": 10000 / dt, "seconds": dt, "workspace_gb": sc.workspace_bytes / 2**30}), flush=True)
    sc.close(); del sc; torch.cuda.empty_cache()
