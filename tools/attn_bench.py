"""Prefill attention kernel alone on a bench-sized batch (Falcon MQA, 71 heads, lengths from the workload histogram)."""
import json, sys
import numpy as np, torch
sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import _lib
import bench
lib = _lib.load()
Lall = bench.synthetic_lengths(10000)
which = sys.argv[1] if len(sys.argv) > 1 else "longest"
L = np.sort(Lall)[::-1][:1180].copy() if which == "longest" else Lall[:1175].copy()
nh, nkv, hd = 71, 1, 64
T = int(L.sum()); LA = 10
qs = np.concatenate([[0], np.cumsum(L)[:-1]]).astype(np.int32)
ks = (qs + np.arange(len(L)) * LA).astype(np.int32)
slots = T + len(L) * LA
qkv_dim = (nh + 2 * nkv) * hd
i32 = lambda x: torch.tensor(np.asarray(x, dtype=np.int32), device="cuda")
a_qs, a_ql, a_ks, a_kl = i32(qs), i32(L), i32(ks), i32(L)
for dt in (torch.bfloat16,):
    qkv = torch.randn(T, qkv_dim, device="cuda").to(dt)
    kc = torch.randn(slots, nkv * hd, device="cuda").to(dt); vc = torch.randn(slots, nkv * hd, device="cuda").to(dt)
    out = torch.zeros(T, nh * hd, device="cuda", dtype=dt)
    for hold in (0, 1, 0, 1):
        lib.b2s_set_option(b"attn_ts", hold)
        ts = []
        for it in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.check(lib.b2s_attention_bf16(_lib.ptr(qkv), _lib.ptr(kc), _lib.ptr(vc), _lib.ptr(out), _lib.ptr(a_qs), _lib.ptr(a_ql),
                                              _lib.ptr(a_ks), _lib.ptr(a_kl), len(L), T, slots, int(L.max()), int(L.max()), nh, nkv, hd,
                                              qkv_dim, 0, hd, 1, _lib.current_stream()), "attn")
            e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        print(json.dumps({"batch": which, "attn_ts": hold, "tokens": T, "ms": round(sorted(ts)[2], 3)}), flush=True)
