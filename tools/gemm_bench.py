"""Micro-benchmark of the tcgen05 GEMM on the Falcon-7B layer shapes (CUDA events, L2 flushed)."""
import json
import sys
import torch
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from llm_interpretation_replication_b200 import _lib


def bench(M, N, K, epi, iters=5):
    lib = _lib.load()
    A = (torch.randn(M, K, device="cuda") * 1.0).to(torch.bfloat16)
    W = (torch.randn(N, K, device="cuda") * 0.02).to(torch.bfloat16)
    out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi in (2, 4) else torch.bfloat16)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    st = _lib.current_stream()
    ts = []
    for i in range(iters + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.b2s_gemm_bf16(_lib.ptr(A), _lib.ptr(W), None, _lib.ptr(out), M, N, K, epi, st), "gemm")
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    # cuBLAS for comparison
    tsb = []
    for i in range(iters + 2):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ref = A @ W.T
        e1.record()
        torch.cuda.synchronize()
        if i >= 2:
            tsb.append(e0.elapsed_time(e1))
    msb = sorted(tsb)[len(tsb) // 2]
    fl = 2.0 * M * N * K
    return dict(M=M, N=N, K=K, epi=epi, ms=ms, tflops=fl / ms / 1e9, cublas_ms=msb, cublas_tflops=fl / msb / 1e9)


if __name__ == "__main__":
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
    res = []
    if len(sys.argv) > 2 and sys.argv[2] == "ncu":     # one launch per shape, for an ncu --set full capture
        lib = _lib.load()
        for (N, K, epi) in ((18176, 4544, 1), (4544, 18176, 2), (4672, 4544, 0)):
            A = (torch.randn(M, K, device="cuda")).to(torch.bfloat16)
            W = (torch.randn(N, K, device="cuda") * 0.02).to(torch.bfloat16)
            out = torch.zeros(M, N, device="cuda", dtype=torch.float32 if epi in (2, 4) else torch.bfloat16)
            for _ in range(2):
                _lib.check(lib.b2s_gemm_bf16(_lib.ptr(A), _lib.ptr(W), None, _lib.ptr(out), M, N, K, epi, _lib.current_stream()), "gemm")
            torch.cuda.synchronize()
        sys.exit(0)
    for (N, K, epi) in ((4672, 4544, 0), (4544, 4544, 2), (18176, 4544, 1), (4544, 18176, 2), (8192, 8192, 0)):
        r = bench(M, N, K, epi)
        print(json.dumps(r), flush=True)
        res.append(r)
