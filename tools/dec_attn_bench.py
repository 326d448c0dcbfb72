"""One-token (look-ahead) attention alone on a bench-sized batch: 1175 prompts, one query row each, Falcon MQA."""
import json, sys
import numpy as np, torch
sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import _lib
import bench
lib = _lib.load()
L = bench.synthetic_lengths(10000)[:1175].copy() + 5        # keys in the cache at look-ahead step 5
nh, nkv, hd = 71, 1, 64
n = len(L)
qs = np.arange(n, dtype=np.int32)
ks = np.concatenate([[0], np.cumsum(L + 5)[:-1]]).astype(np.int32)
slots = int((L + 5).sum()) + 300
qkv_dim = (nh + 2 * nkv) * hd
i32 = lambda x: torch.tensor(np.asarray(x, dtype=np.int32), device="cuda")
a_qs, a_ql, a_ks, a_kl = i32(qs), i32(np.ones(n)), i32(ks), i32(L)
dt = torch.bfloat16
qkv = torch.randn(n, qkv_dim, device="cuda").to(dt)
kc = torch.randn(slots, nkv * hd, device="cuda").to(dt); vc = torch.randn(slots, nkv * hd, device="cuda").to(dt)
out = torch.zeros(n, nh * hd, device="cuda", dtype=dt)
for ts in (0, 1, 0, 1):
    lib.b2s_set_option(b"attn_ts", ts)
    t = []
    for it in range(8):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.b2s_attention_bf16(_lib.ptr(qkv), _lib.ptr(kc), _lib.ptr(vc), _lib.ptr(out), _lib.ptr(a_qs), _lib.ptr(a_ql),
                                          _lib.ptr(a_ks), _lib.ptr(a_kl), n, n, slots, 1, int(L.max()), nh, nkv, hd,
                                          qkv_dim, 0, hd, 1, _lib.current_stream()), "attn")
        e1.record(); torch.cuda.synchronize()
        t.append(e0.elapsed_time(e1))
    print(json.dumps({"attn_ts": ts, "prompts": n, "max_kv": int(L.max()), "us": round(sorted(t)[3] * 1000, 1)}), flush=True)
lib.b2s_set_option(b"attn_ts", 1)
