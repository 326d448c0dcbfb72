"""What LLM.int8 (the reference's USE_8BIT configuration, emulated: MODE_INT8_EMU) does to relative_prob at full depth:
Falcon-7B, 32 layers, random-init, bench prompts, against the FP16 mode (itself within 9e-4 of the fp32 check mode,
profiles/r2_parity_histogram.json).  Also with the outlier threshold off (int8_threshold_milli = 0: every column quantised)."""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
from llm_interpretation_replication_b200 import _lib, shapes, weights  # noqa: E402
from llm_interpretation_replication_b200.scorer import Scorer  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
shp = shapes.falcon_7b()
ids, cand, L = bench.synthetic_prompts(shp, n, seed=5)
W = weights.init_weights_device(shp, seed=0, device="cuda:0", weight_dtype=torch.bfloat16)
lib = _lib.load()
rows = {}
for name, mode, thr in (("fp16", shapes.MODE_FP16, None), ("int8_emu", shapes.MODE_INT8_EMU, 6000), ("int8_emu_no_outlier_columns", shapes.MODE_INT8_EMU, 0)):
    if thr is not None:
        lib.b2s_set_option(b"int8_threshold_milli", thr)
    sc = Scorer(shp, W, mode=mode, device="cuda:0", max_tokens=131072, max_seqs=2048)
    rows[name], _ = sc.score_ids(ids, cand, topk=5, nonfinite="nan")
    sc.close()
    del sc
    torch.cuda.empty_cache()
lib.b2s_set_option(b"int8_threshold_milli", 6000)
ref = rows["fp16"]
for name in ("int8_emu", "int8_emu_no_outlier_columns"):
    r = rows[name]
    same = ref[:, 6] == r[:, 6]
    d = np.abs(r[same, 2] - ref[same, 2])
    print(json.dumps({"model": "falcon-7b, 32 layers, random-init", "prompts": n, "mode": name, "vs": "fp16 mode",
                      "position_disagreements": int((~same).sum()),
                      "abs_delta_relative_prob": {"mean": float(d.mean()), "median": float(np.median(d)), "p95": float(np.percentile(d, 95)),
                                                  "max": float(d.max())},
                      "within_2e-3": float((d <= 2e-3).mean()), "within_1e-2": float((d <= 1e-2).mean())}), flush=True)
