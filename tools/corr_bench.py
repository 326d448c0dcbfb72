"""N3 timing: 1000 prompt resamples of the 12-model x 60-prompt golden table and of a 18-model x 2000-prompt table,
GPU (b2s_corr_bootstrap through correlations.py) next to the reference loop's pandas calls on the host."""
import json, os, sys, time
import numpy as np, pandas as pd, torch
sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import correlations as Cn

def table(P, M, seed=0):
    g = np.random.RandomState(seed)
    lat = g.rand(P)
    rows = [dict(prompt=f"p{p}", model=f"m{m}", relative_prob=float(np.clip(lat[p] + 0.2 * g.randn(), 0, 1))) for m in range(M) for p in range(P)]
    return pd.DataFrame(rows)

for name, df in (("golden 12x60", pd.read_csv("tests/golden/corr_input.csv")), ("18x2000", table(2000, 18))):
    for method in ("pearson", "spearman"):
        Cn.calculate_model_correlations(df, method, 1000, 1000)      # warm-up (module load, allocator)
        torch.cuda.synchronize(); t0 = time.time()
        Cn.calculate_model_correlations(df, method, 1000, 1000)
        torch.cuda.synchronize(); t_gpu = time.time() - t0
        piv = df.pivot_table(index="prompt", columns="model", values="relative_prob")
        g = np.random.RandomState(42); P = len(piv)
        t0 = time.time(); nb = 50
        for _ in range(nb):
            piv.iloc[g.choice(P, P)].corr(method=method)
        t_cpu = (time.time() - t0) / nb * 1000
        print(json.dumps({"table": name, "method": method, "gpu_total_s_1000_resamples": round(t_gpu, 4),
                          "pandas_s_1000_resamples_extrapolated": round(t_cpu, 3)}), flush=True)
