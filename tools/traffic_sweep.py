"""One launch per (layer shape, tiling, rasterisation) for an ncu metrics pass (DRAM / L2 bytes per launch), or --time
for a sustained-throughput pass of the same grid on the same box.  fp16 operands as in the default bench mode.
Prints one JSON line per launch in launch order so the ncu CSV rows can be joined by index."""
import json, sys, time
import torch
sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import _lib

M = int(sys.argv[1]) if len(sys.argv) > 1 else 111104
timed = "--time" in sys.argv
secs = 1.0
lib = _lib.load(); st = _lib.current_stream()
shapes = {"up": (18176, 4544, 1), "down": (4544, 18176, 2), "dense": (4544, 4544, 2), "qkv": (4672, 4544, 0)}
if "--wavesync" in sys.argv:
    grid = [(cg, 0, 0, ws) for cg in (3, 2) for ws in (0, 1, 2, 4, 8)] + [(1, 0, 0, 0)]
    shapes = {k: shapes[k] for k in ("down", "up")}
elif "--pairs" in sys.argv:
    grid = [(cg, 0, 0, ps) for cg in (2, 3) for ps in (1, 3, 9, 21, 27, 35)] + [(1, 0, 0, 1)]
    shapes = {k: shapes[k] for k in ("down", "up")}
elif "--cg3" in sys.argv:
    grid = [(3, gm, 0, 0) for gm in (4, 8, 12, 16, 24, 32, 48, 64, 128)] + [(1, 0, 0, 0), (2, 0, 0, 0)]
    shapes = {k: shapes[k] for k in ("down", "up")}
elif "--hints" in sys.argv:
    grid = [(0, gm, 1, h) for gm in (0, 8) for h in (0, 4, 5, 6, 7, 1)]
else:
    grid = [(cg, gm, serp, 0) for cg in (0,) for gm in (0, 8, 16, 32, 64, 128) for serp in (0, 1)]
    grid += [(3, 0, 0, 0), (3, 0, 1, 0), (3, 64, 1, 0), (2, 0, 1, 0)]
for name, (N, K, epi) in shapes.items():
    A = torch.randn(M, K, device="cuda").to(torch.float16)
    W = (torch.randn(N, K, device="cuda") * 0.02).to(torch.float16)
    out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi == 2 else torch.float16)
    alg = 2 * (M * K + N * K) + (8 if epi == 2 else 2) * M * N
    for cg, gm, serp, hint in grid:
        def f():
            lib.b2s_set_option(b"gemm_cta_group", cg); lib.b2s_set_option(b"gemm_group_m", gm); lib.b2s_set_option(b"gemm_serpentine", serp); lib.b2s_set_option(b"gemm_l2_hint", 0 if ("--pairs" in sys.argv or "--wavesync" in sys.argv) else hint); lib.b2s_set_option(b"gemm_pair_stride", hint if "--pairs" in sys.argv else 1); lib.b2s_set_option(b"gemm_wave_sync", hint if "--wavesync" in sys.argv else 0)
            _lib.check(lib.b2s_gemm_f16(_lib.ptr(A), _lib.ptr(W), None, _lib.ptr(out), M, N, K, epi, st), "gemm")
        rec = {"shape": name, "cg": cg, "group_m": gm, "serp": serp, "hint": hint, "algorithmic_bytes": alg}
        if timed:
            f(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.time(); n = 0; e0.record()
            while time.time() - t0 < secs:
                f(); n += 1
                if n % 8 == 0: torch.cuda.synchronize()
            e1.record(); torch.cuda.synchronize()
            rec["tflops"] = round(n * 2.0 * M * N * K / e0.elapsed_time(e1) / 1e9, 1)
        else:
            f(); torch.cuda.synchronize()
        print(json.dumps(rec), flush=True)
    del A, W, out
