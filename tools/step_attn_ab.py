"""A/B of the streaming one-token attention kernel (`attn_step` option) on decoder-only shapes whose look-ahead steps do not
take the tensor-core one-token kernel: Mistral-7B (head_dim 128, grouped-query 32/8), StableLM-alpha-7B (head_dim 128, 48 heads),
OPT-1.3B (head_dim 64, multi-head), 8 layers each, 2000 perturbation-length prompts, top-5 / 10-position look-ahead."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
from llm_interpretation_replication_b200 import _lib, shapes, weights  # noqa: E402
from llm_interpretation_replication_b200.scorer import Scorer  # noqa: E402

lib = _lib.load()
for maker in ("mistral_7b", "stablelm_alpha_7b", "opt_iml_1_3b"):
    shp = getattr(shapes, maker)().with_layers(8)
    ids, cand, L = bench.synthetic_prompts(shp, 2000, seed=2)
    W = weights.init_weights_device(shp, seed=1, device="cuda:0", weight_dtype=torch.float16)
    sc = Scorer(shp, W, mode=shapes.MODE_FP16, device="cuda:0", max_tokens=262144, max_seqs=4096)
    res = {}
    rows = {}
    for opt in (1, 0, 1, 0):
        lib.b2s_set_option(b"attn_step", opt)
        sc.score_ids(ids[:300], cand[:300], topk=5)
        torch.cuda.synchronize()
        t0 = time.time()
        rows[opt], _ = sc.score_ids(ids, cand, topk=5, nonfinite="nan")
        torch.cuda.synchronize()
        res.setdefault(opt, []).append(time.time() - t0)
    lib.b2s_set_option(b"attn_step", 1)
    print(json.dumps({"shape": shp.name, "prompts": len(ids), "seconds_streaming": min(res[1]), "seconds_varlen_or_tc": min(res[0]),
                      "speedup": min(res[0]) / min(res[1]),
                      "max_abs_delta_rel_prob_between_the_two": float(np.nanmax(np.abs(rows[1][:, 2] - rows[0][:, 2])))}), flush=True)
    sc.close()
    del sc, W
    torch.cuda.empty_cache()
