"""Sustained-throughput sweep of rasterisation group size / L2 hints / tiling per layer shape (same box, power-capped)."""
import json, subprocess, sys, time
import torch
sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import _lib
sys.path.insert(0, "tools")
from sustained_gemm import sample_start, sample_stop

M = int(sys.argv[1]) if len(sys.argv) > 1 else 111104
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 1.5
lib = _lib.load(); st = _lib.current_stream()
shapes = {"up": (18176, 4544, 1), "down": (4544, 18176, 2), "dense": (4544, 4544, 2), "qkv": (4672, 4544, 0)}
epi_override = int(sys.argv[4]) if len(sys.argv) > 4 else -1
which = sys.argv[3].split(",") if len(sys.argv) > 3 else ["down", "up", "dense", "qkv"]
for name in which:
    N, K, epi = shapes[name]
    if epi_override >= 0: epi = epi_override
    A = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    W = (torch.randn(N, K, device="cuda") * 0.02).to(torch.bfloat16)
    out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi == 2 else torch.bfloat16)
    def run(tag, fn):
        fn(); torch.cuda.synchronize()
        p = sample_start(); t0 = time.time(); n = 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        while time.time() - t0 < secs:
            fn(); n += 1
            if n % 8 == 0: torch.cuda.synchronize()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1); clk, pw = sample_stop(p)
        print(json.dumps({"shape": name, "cfg": tag, "tflops": round(n * 2.0 * M * N * K / ms / 1e9, 1), "sm_mhz": clk, "power_w": pw}), flush=True)
    run("cublas", lambda: A @ W.T)
    for cg in (1, 2, 3):
        for gm in (0,):
            for hint in (0,):
                def f(cg=cg, gm=gm, hint=hint):
                    lib.b2s_set_option(b"gemm_cta_group", cg); lib.b2s_set_option(b"gemm_group_m", gm); lib.b2s_set_option(b"gemm_l2_hint", hint)
                    _lib.check(lib.b2s_gemm_bf16(_lib.ptr(A), _lib.ptr(W), None, _lib.ptr(out), M, N, K, epi, st), "gemm")
                run(f"cg{cg}_gm{gm}_hint{hint}", f)
    run("cublas_again", lambda: A @ W.T)
    del A, W, out
