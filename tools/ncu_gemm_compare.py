"""One launch each of cuBLAS (torch.matmul), our single-CTA and our CTA-pair GEMM on the up / down shapes, for an
ncu metrics pass (DRAM bytes, L2 bytes, executed instructions, tensor-pipe activity)."""
import sys
import torch
sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import _lib

M = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
lib = _lib.load(); st = _lib.current_stream()
for (N, K, epi) in ((4544, 18176, 0), (18176, 4544, 0)):     # plain bf16 store epilogue for all three: like for like
    A = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    W = (torch.randn(N, K, device="cuda") * 0.02).to(torch.bfloat16)
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    for rep in range(2):
        ref = A @ W.T
        for cg in (1, 2, 4):
            lib.b2s_set_option(b"gemm_cta_group", cg)
            _lib.check(lib.b2s_gemm_bf16(_lib.ptr(A), _lib.ptr(W), None, _lib.ptr(out), M, N, K, epi, st), "gemm")
        torch.cuda.synchronize()
    del A, W, out, ref
