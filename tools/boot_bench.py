"""Bootstrap kernels at the BASELINE config-5 size (10,000 x 100, seed 42) and at a scaled size where the
resample kernel's achieved GB/s is meaningful, next to the reference functions' NumPy loops
(oracle restatement, 1 host core).  Writes one JSON object."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import bootstrap as B


def ev():
    return torch.cuda.Event(enable_timing=True)


def main():
    rs = np.random.RandomState(0)
    base = np.abs(rs.normal(0.3, 0.15, 100)); inst = np.abs(rs.normal(0.35, 0.2, 100))
    out = {}
    # C5 through the public wrapper (seed + index stream + resample + sort/percentile + D2H of the scalars)
    B.bootstrap_mae_difference(base, inst)
    torch.cuda.synchronize()
    t = time.time()
    for _ in range(20):
        B.bootstrap_mae_difference(base, inst)
    torch.cuda.synchronize()
    out["c5_gpu_ms_per_call"] = (time.time() - t) / 20 * 1e3
    from oracle import bootstrap_oracle
    t = time.time()
    bootstrap_oracle.bootstrap_mae_difference(base, inst)
    out["c5_cpu_reference_ms_per_call"] = (time.time() - t) * 1e3
    out["c5_resamples_per_s_gpu"] = 10000 / (out["c5_gpu_ms_per_call"] / 1e3)
    out["c5_resamples_per_s_cpu"] = 10000 / (out["c5_cpu_reference_ms_per_call"] / 1e3)
    # kernel-level: MT19937 stream and the resample kernel
    st = B.NumpyLegacyStream(42)
    e0, e1 = ev(), ev()
    e0.record(); idx = st.choice(100, 1_000_000); e1.record(); torch.cuda.synchronize()
    out["mt19937_choice_1M_draws_ms"] = e0.elapsed_time(e1)
    vals = np.stack([base, inst])
    n, Bn = 100, (1_000_000 if "--small" in sys.argv else 10_000_000)    # --small: short enough for an ncu --set full replay
    idx = torch.randint(0, n, (Bn * n,), dtype=torch.int32, device="cuda")     # index table of the scaled size (4 GB)
    boot = B.resample(vals, 1, Bn, idx)
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        e0, e1 = ev(), ev()
        e0.record(); B.resample(vals, 1, Bn, idx); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[1]
    byts = Bn * (4 * n + 24) + 16 * n
    out["resample_scaled"] = {"B": Bn, "n": n, "ms": ms, "algorithmic_bytes": byts, "achieved_gbs": byts / ms / 1e6}
    try:
        peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    out["resample_scaled"]["frac_of_measured_hbm"] = out["resample_scaled"]["achieved_gbs"] / peak
    e0, e1 = ev(), ev()
    e0.record(); B.resample(vals, 1, Bn, None, philox_seed=7); e1.record(); torch.cuda.synchronize()
    out["resample_philox_scaled_ms"] = e0.elapsed_time(e1)
    out["resample_philox_gdraws_per_s"] = Bn * n / out["resample_philox_scaled_ms"] / 1e6
    print(json.dumps(out))


if __name__ == "__main__":
    main()
