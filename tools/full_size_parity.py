"""Full-size parity probe: Falcon-7B shape, all 32 layers, bf16 tensor-core path on the GPU vs the
fp32 HF forward on the host over the SAME (bf16-representable) weights, position-0 relative probability.
Prints the |delta rel_prob| distribution and the pass rate against the 2e-3 gate.  Not part of pytest (minutes of
CPU time, 28 GB of host memory)."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import shapes, weights
from llm_interpretation_replication_b200.scorer import Scorer
import bench


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 12
    n_layer = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    modes = {"bf16": shapes.MODE_BF16_FAST, "fp16": shapes.MODE_FP16}
    which = sys.argv[3].split(",") if len(sys.argv) > 3 else ["bf16", "fp16"]
    shp = shapes.falcon_7b()
    if n_layer != 32:
        shp = shp.with_layers(n_layer)
    dev = torch.device("cuda:0")
    W = weights.init_weights_device(shp, seed=0, device=dev, weight_dtype=torch.bfloat16, randomize_affine=True)
    ids, cand, L = bench.synthetic_prompts(shp, 10000)
    pick = np.linspace(0, 9999, n).astype(int)
    my_ids = [ids[i].tolist() for i in pick]
    my_cand = cand[pick]
    all_rows = {}
    for name in which:
        sc = Scorer(shp, W, mode=modes[name], device=dev, max_tokens=8192, max_seqs=64, max_lookahead=1)
        all_rows[name], _ = sc.score_ids(my_ids, my_cand, topk=5, max_lookahead=1)
        sc.close()
    t0 = time.time()
    Wc = {k: v.detach().to("cpu", torch.float32) for k, v in W.items() if k not in ("rope_cos", "rope_sin")}
    Wc["lm_head"] = Wc["wte"]
    model = weights.build_hf_model(shp, Wc)
    torch.set_num_threads(torch.get_num_threads())
    print(f"host model ready in {time.time() - t0:.1f}s", flush=True)
    errs = {name: [] for name in which}
    rels = []
    for k, x in enumerate(my_ids):
        with torch.no_grad():
            lo = model(torch.tensor([x])).logits[0, -1].float()
        p = torch.softmax(lo, -1)
        a, b = int(my_cand[k][0]), int(my_cand[k][1])
        rel = (p[a] / (p[a] + p[b])).item()
        rels.append(rel)
        msg = f"prompt {pick[k]} L={len(x)} rel_ref={rel:.6f}"
        for name in which:
            errs[name].append(abs(rel - float(all_rows[name][k][2])))
            msg += f" | {name}: rel={all_rows[name][k][2]:.6f} d={errs[name][-1]:.2e}"
        print(msg, flush=True)
    for name in which:
        e = np.array(errs[name])
        print(json.dumps({"mode": name, "n": n, "n_layer": n_layer, "max": float(e.max()), "mean": float(e.mean()),
                          "pass_2e-3": float((e <= 2e-3).mean()), "rel_spread": float(np.std(rels))}))


if __name__ == "__main__":
    main()
