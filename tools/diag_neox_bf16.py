"""Noise or bug?  StableLM-alpha-7B widths, 2 layers: relative_prob from the GPU (bf16 and fp16 modes, epilogue through
TMA staging and through direct stores) beside the fp32 oracle and the bf16-rounding oracle."""
import sys, os, json
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import gpu_util as U
from llm_interpretation_replication_b200 import shapes, weights, _lib
from oracle import decoder_ref

shp = shapes.stablelm_alpha_7b().with_layers(2)
W = weights.init_weights(shp, seed=5)
g = np.random.RandomState(2)
ids = [g.randint(32, shp.vocab - 32, size=L).tolist() for L in (18, 84, 40)]
def orc(rnd):
    out = []
    for x in ids:
        lo = decoder_ref.last_logits(shp, W, torch.tensor(x), round_act=rnd)
        p = torch.softmax(lo, -1)
        out.append((p[500] / (p[500] + p[600])).item())
    return np.array(out)
o32 = orc(lambda t: t)
ob = orc(decoder_ref.bf16_round)
oh = orc(lambda t: t.to(torch.float16).to(t.dtype))
print(json.dumps({"oracle_fp32": o32.tolist(), "oracle_bf16_minus_fp32": (ob - o32).tolist(), "oracle_fp16_minus_fp32": (oh - o32).tolist()}))
lib = _lib.load()
for mode_name in ("MODE_BF16_FAST", "MODE_FP16", "MODE_FP32_CHECK"):
    for tma in (1, 0):
        lib.b2s_set_option(b"gemm_tma_store", tma)
        res = []
        for rep in range(2):
            sc = U.make_scorer(shp, W, getattr(shapes, mode_name), max_tokens=1024, max_seqs=8, max_lookahead=2)
            rows, _ = sc.score_ids(ids, np.array([[500, 600]]), topk=5, max_lookahead=2)
            res.append(np.array([float(r[2]) for r in rows]))
            sc.close()
        print(json.dumps({"mode": mode_name, "tma_store": tma, "gpu_minus_fp32": (res[0] - o32).tolist(),
                          "gpu_minus_bf16_oracle": (res[0] - ob).tolist(), "repeat_identical": bool((res[0] == res[1]).all())}))
lib.b2s_set_option(b"gemm_tma_store", 1)
