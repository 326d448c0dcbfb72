"""Micro-benchmark of the tcgen05 GEMM on the Falcon-7B layer shapes (CUDA events, L2 flushed).
The variants (one CTA per tile, CTA pairs, cuBLAS via torch.matmul) are interleaved round by round so that clock /
power drift during the run hits all of them alike; medians over the rounds are reported."""
import json
import sys
import torch
sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import _lib

SHAPES = ((4672, 4544, 0), (4544, 4544, 2), (18176, 4544, 1), (4544, 18176, 2))


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
    lib = _lib.load()
    st = _lib.current_stream()
    if len(sys.argv) > 2 and sys.argv[2] == "ncu":     # one launch per shape and tiling, for an ncu --set full capture
        for (N, K, epi) in SHAPES:
            A = torch.randn(M, K, device="cuda").to(torch.bfloat16)
            W = (torch.randn(N, K, device="cuda") * 0.02).to(torch.bfloat16)
            out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi in (2, 4) else torch.bfloat16)
            for cg in (1, 2):
                _lib.check(lib.b2s_set_option(b"gemm_cta_group", cg), "opt")
                _lib.check(lib.b2s_gemm_bf16(_lib.ptr(A), _lib.ptr(W), None, _lib.ptr(out), M, N, K, epi, st), "gemm")
            torch.cuda.synchronize()
        return
    rounds = 7
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for (N, K, epi) in SHAPES:
        A = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        W = (torch.randn(N, K, device="cuda") * 0.02).to(torch.bfloat16)
        out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi in (2, 4) else torch.bfloat16)
        t = {"cta1": [], "cta_pair": [], "cublas": []}

        def timed(fn):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1)

        def ours(cg):
            _lib.check(lib.b2s_set_option(b"gemm_cta_group", cg), "opt")
            _lib.check(lib.b2s_gemm_bf16(_lib.ptr(A), _lib.ptr(W), None, _lib.ptr(out), M, N, K, epi, st), "gemm")
        for r in range(rounds + 1):
            a = timed(lambda: ours(1)); b = timed(lambda: ours(2)); c = timed(lambda: A @ W.T)
            if r:
                t["cta1"].append(a); t["cta_pair"].append(b); t["cublas"].append(c)
        fl = 2.0 * M * N * K
        med = {k: sorted(v)[len(v) // 2] for k, v in t.items()}
        print(json.dumps({"M": M, "N": N, "K": K, "epi": epi, **{k + "_ms": round(v, 4) for k, v in med.items()},
                          **{k + "_tflops": round(fl / v / 1e9, 1) for k, v in med.items()}}), flush=True)
    _lib.check(lib.b2s_set_option(b"gemm_cta_group", 0), "opt")


if __name__ == "__main__":
    main()
