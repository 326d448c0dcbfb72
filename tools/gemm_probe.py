"""One launch per (layer shape, knob setting) for an ncu metrics pass that asks WHERE the tensor pipe idles:
    ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum \
        --clock-control none -k regex:gemm_tc --csv --log-file gpurun_out/gemm_probe.csv python tools/gemm_probe.py
Prints one JSON line per launch, in launch order, so the CSV rows can be joined by index.  fp16 operands, M = 131072."""
import json
import sys

import torch

sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import _lib

M = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
lib = _lib.load()
st = _lib.current_stream()
DEFAULTS = {"gemm_cta_group": 0, "gemm_wave_sync": 4, "gemm_band_n": -1, "gemm_prefetch": 0, "gemm_group_m": 0, "gemm_tma_store": 1}
shapes = {"up": (18176, 4544), "dense": (4544, 4544), "qkv": (4672, 4544), "down": (4544, 18176)}
# (label, epilogue or None = the layer's own, knob overrides)
configs = [("default", None, {}), ("no_meeting", None, {"gemm_wave_sync": 0}), ("meeting_every_8", None, {"gemm_wave_sync": 8}),
           ("plain_store_epilogue", 0, {}), ("fp32_store_epilogue", 4, {}), ("pairs_512x256", None, {"gemm_cta_group": 3}),
           ("single_cta", None, {"gemm_cta_group": 1}), ("row_groups_only", None, {"gemm_band_n": 0}),
           ("group_m_8", None, {"gemm_group_m": 8}), ("group_m_64", None, {"gemm_group_m": 64})]
own_epi = {"up": 1, "dense": 2, "qkv": 0, "down": 2}
for name, (N, K) in shapes.items():
    A = torch.randn(M, K, device="cuda").to(torch.float16)
    W = (torch.randn(N, K, device="cuda") * 0.02).to(torch.float16)
    o16 = torch.zeros(M, N, device="cuda", dtype=torch.float16)
    o32 = torch.zeros(M, N, device="cuda", dtype=torch.float32)
    for label, epi, knobs in configs:
        e = own_epi[name] if epi is None else epi
        for k, v in {**DEFAULTS, **knobs}.items():
            _lib.check(lib.b2s_set_option(k.encode(), v), k)
        out = o32 if e in (2, 4) else o16
        _lib.check(lib.b2s_gemm_f16(_lib.ptr(A), _lib.ptr(W), None, _lib.ptr(out), M, N, K, e, st), "gemm")
        torch.cuda.synchronize()
        print(json.dumps({"shape": name, "config": label, "epilogue": e, "M": M, "N": N, "K": K, "flops": 2.0 * M * N * K}), flush=True)
    del A, W, o16, o32
    torch.cuda.empty_cache()
for k, v in DEFAULTS.items():
    lib.b2s_set_option(k.encode(), v)
