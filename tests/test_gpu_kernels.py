"""GPU parity of the building-block kernels, called through the C-ABI."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(params=[1, 2, 4], ids=["cta1", "cta_pair", "cta_pair_big"])
def cta_group(request):
    """Run the tensor-core GEMM tests on every tiling: one CTA per 128x256 tile, CTA pairs (cta_group::2) on 256x256
    tiles, and CTA pairs on 512x256 tiles (4 = forced also for small M)."""
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    _lib.check(lib.b2s_set_option(b"gemm_cta_group", request.param), "set_option")
    yield request.param
    _lib.check(lib.b2s_set_option(b"gemm_cta_group", 0), "set_option")   # library default


def _rand(shape, scale=1.0, seed=0, dtype=torch.bfloat16):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(shape, generator=g, device="cuda") * scale).to(dtype)


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (128, 256, 512), (256, 512, 4544), (300, 520, 192),
                                   (77, 4672, 4544), (1, 264, 72), (1000, 4544, 1024), (513, 18176, 4544)])
def test_gemm_tc_store_f32(M, N, K, cta_group):
    import gpu_util as U
    A, W = _rand((M, K), 1.0, 1), _rand((N, K), 0.05, 2)
    ref = A.float() @ W.float().T
    out = U.gemm_bf16(A, W, None, 4)
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    assert err <= 2e-3 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("serp", [1, 2], ids=["by_column", "by_wave"])
@pytest.mark.parametrize("epi", [4, 2])
def test_gemm_tc_serpentine_rasterisation(epi, serp, cta_group):
    """Serpentine order (odd groups walk column tiles backwards; K walked backwards by odd column tiles or odd waves;
    K tail of 8): several groups and several waves per CTA, same result as the plain order."""
    import gpu_util as U
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    M, N, K = 4200, 2600, 520
    A, W = _rand((M, K), 1.0, 11), _rand((N, K), 0.05, 12)
    ref = A.float() @ W.float().T
    resid = _rand((M, N), 1.0, 13, torch.float32)
    try:
        _lib.check(lib.b2s_set_option(b"gemm_serpentine", serp), "set_option")
        _lib.check(lib.b2s_set_option(b"gemm_group_m", 4), "set_option")
        out = U.gemm_bf16(A, W, None, epi, out=resid.clone() if epi == 2 else None)
        torch.cuda.synchronize()
    finally:
        _lib.check(lib.b2s_set_option(b"gemm_serpentine", 0), "set_option")
        _lib.check(lib.b2s_set_option(b"gemm_group_m", 0), "set_option")
    if epi == 2:
        ref = ref + resid
    err = (out - ref).abs().max().item()
    assert err <= 2e-3 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("epi", [0, 1, 2])
@pytest.mark.parametrize("use_bias", [False, True])
def test_gemm_tc_epilogues(epi, use_bias, cta_group):
    import gpu_util as U
    M, N, K = 391, 1160, 832
    A, W = _rand((M, K), 1.0, 3), _rand((N, K), 0.05, 4)
    bias = _rand((N,), 0.5, 5, torch.float32) if use_bias else None
    ref = A.float() @ W.float().T
    if use_bias:
        ref = ref + bias
    if epi == 1:
        ref = torch.nn.functional.gelu(ref)
    if epi == 2:
        resid = _rand((M, N), 1.0, 6, torch.float32)
        out = U.gemm_bf16(A, W, bias, 2, out=resid.clone())
        ref = ref + resid
        tol = 2e-3
    else:
        out = U.gemm_bf16(A, W, bias, epi).float()
        tol = 2.0 ** -7   # one bf16 ulp relative
    torch.cuda.synchronize()
    err = ((out - ref).abs() / (ref.abs() + 1.0)).max().item()
    assert err <= tol, err


@pytest.mark.parametrize("M,N,K,epi", [(300, 520, 192, 4), (513, 4672, 4544, 0), (391, 1160, 832, 1), (200, 264, 4544, 2)])
def test_gemm_tc_fp16_operands(M, N, K, epi, cta_group):
    """FP16 mode: both operands fp16 (tcgen05 kind::f16 with a_format = b_format = F16)."""
    import gpu_util as U
    A, W = _rand((M, K), 1.0, 21, torch.float16), _rand((N, K), 0.05, 22, torch.float16)
    ref = A.float() @ W.float().T
    if epi == 1:
        ref = torch.nn.functional.gelu(ref)
    if epi == 2:
        resid = _rand((M, N), 1.0, 23, torch.float32)
        out = U.gemm_f16(A, W, None, 2, out=resid.clone())
        ref = ref + resid
    else:
        out = U.gemm_f16(A, W, None, epi).float()
    torch.cuda.synchronize()
    tol = 2e-3 if epi in (2, 4) else 2.0 ** -10      # fp16 output: one ulp relative
    err = ((out - ref).abs() / (ref.abs() + 1.0)).max().item()
    assert err <= tol, err


def test_gemm_tc_rejects_bad_shapes():
    import gpu_util as U
    from llm_interpretation_replication_b200 import _lib
    A, W = _rand((8, 20), 1, 1), _rand((16, 20), 1, 2)   # K not a multiple of 8
    with pytest.raises(_lib.B2SError):
        U.gemm_bf16(A, W, None, 4)


@pytest.mark.parametrize("M,N,K", [(64, 64, 16), (100, 70, 33), (257, 513, 129)])
@pytest.mark.parametrize("epi", [4, 1, 2])
def test_gemm_f32(M, N, K, epi):
    import gpu_util as U
    A, W = _rand((M, K), 1.0, 1, torch.float32), _rand((N, K), 0.1, 2, torch.float32)
    bias = _rand((N,), 0.5, 3, torch.float32)
    ref = (A.double() @ W.double().T + bias.double())
    if epi == 1:
        ref = torch.nn.functional.gelu(ref)
    out0 = _rand((M, N), 1.0, 4, torch.float32)
    if epi == 2:
        ref = ref + out0.double()
    out = U.gemm_f32(A, W, bias, epi, out=out0.clone())
    torch.cuda.synchronize()
    assert (out.double() - ref).abs().max().item() <= 1e-4


@pytest.mark.parametrize("M,V,d", [(5, 1000, 192), (128, 65024, 256), (200, 50432, 512), (1, 320, 64)])
def test_lm_head_fused(M, V, d, cta_group):
    import gpu_util as U
    h, W = _rand((M, d), 1.0, 7), _rand((V, d), 0.05, 8)
    g = torch.Generator(device="cuda").manual_seed(9)
    cand = torch.randint(0, V, (M, 2), generator=g, device="cuda", dtype=torch.int32)
    logits = h.float() @ W.float().T
    cand[0, 0] = int(logits[0].argmax())       # a candidate that is the arg max
    cand[-1, 1] = V - 1                        # last column of the (ragged) last tile
    stats = U.lm_head(h, W, cand)
    torch.cuda.synchronize()
    lse = torch.logsumexp(logits, dim=-1)
    top5 = torch.topk(logits, 5, dim=-1).values
    assert (stats[:, 0] - lse).abs().max().item() <= 2e-3
    assert (stats[:, 1:6] - top5).abs().max().item() <= 2e-3
    arg = stats.view(torch.int32)[:, 15].long()
    picked = logits.gather(1, arg[:, None])[:, 0]
    assert (picked - logits.max(dim=-1).values).abs().max().item() <= 2e-3   # arg max up to fp32 summation order
    ca = logits.gather(1, cand.long())
    assert (stats[:, 7:9] - ca).abs().max().item() <= 2e-3


def _attn_ref(q, k, v, lens, kv_extra):
    """fp32 torch reference: per sequence causal softmax(q k^T / sqrt(d)) v with kv_extra earlier keys."""
    outs = []
    for (qs, ql, ks, kl) in lens:
        Q = q[qs:qs + ql].float()                # [ql, nh, hd]
        K = k[ks:ks + kl].float()                # [kl, nkv, hd]
        V = v[ks:ks + kl].float()
        rep = Q.shape[1] // K.shape[1]
        K = K.repeat_interleave(rep, dim=1); V = V.repeat_interleave(rep, dim=1)
        s = torch.einsum("qhd,khd->hqk", Q, K) / (Q.shape[-1] ** 0.5)
        qi = torch.arange(ql, device=q.device)[:, None]; kj = torch.arange(kl, device=q.device)[None, :]
        s = s.masked_fill((kj > (kl - ql) + qi)[None], float("-inf"))
        outs.append(torch.einsum("hqk,khd->qhd", torch.softmax(s, -1), V))
    return outs


def test_attention_simt_kv_longer_than_shared_memory():
    """head_dim 128 with 240 keys (a 190-token prompt + a 50-token completion): K/V no longer fit in shared memory
    as fp32, the CUDA-core kernel reads them through L2 instead (prefill rows and a one-token row)."""
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    nh, nkv, hd = 4, 2, 128
    qlens, extra = [200, 1, 37], [40, 239, 0]
    q_start, kv_start, t, c = [], [], 0, 0
    for ql, ex in zip(qlens, extra):
        q_start.append(t); kv_start.append(c); t += ql; c += ql + ex + 7
    T, slots = t, c + 16
    qkv_dim = (nh + 2 * nkv) * hd
    qkv = _rand((T, qkv_dim), 1.0, 21)
    kc, vc = _rand((slots, nkv * hd), 1.0, 22), _rand((slots, nkv * hd), 1.0, 23)
    out = torch.zeros((T, nh * hd), dtype=torch.bfloat16, device="cuda")
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device="cuda")
    kvl = [a + b for a, b in zip(qlens, extra)]
    a_qs, a_ql, a_ks, a_kl = i32(q_start), i32(qlens), i32(kv_start), i32(kvl)     # keep alive across the launch
    _lib.check(lib.b2s_attention_bf16(_lib.ptr(qkv), _lib.ptr(kc), _lib.ptr(vc), _lib.ptr(out), _lib.ptr(a_qs),
                                      _lib.ptr(a_ql), _lib.ptr(a_ks), _lib.ptr(a_kl), len(qlens), T, slots,
                                      max(qlens), max(kvl), nh, nkv, hd, qkv_dim, 0, hd, 0, _lib.current_stream()), "attention")
    torch.cuda.synchronize()
    q = qkv[:, :nh * hd].view(T, nh, hd)
    ref = _attn_ref(q, kc.view(slots, nkv, hd), vc.view(slots, nkv, hd), list(zip(q_start, qlens, kv_start, kvl)), extra)
    for (qs, ql), r in zip(zip(q_start, qlens), ref):
        err = (out[qs:qs + ql].view(ql, nh, hd).float() - r).abs().max().item()
        assert err <= 1.5e-2, (ql, err)


@pytest.mark.parametrize("use_tc", [0, 1, 2], ids=["simt", "tc_p_in_tmem", "tc_p_in_smem"])
@pytest.mark.parametrize("nh,nkv", [(71, 1), (4, 2), (3, 3)])
@pytest.mark.parametrize("phase", ["prefill", "one_token", "one_token_short"])
def test_attention_bf16(use_tc, nh, nkv, phase):
    """Ragged prefill batch (lengths 1..256, incl. exactly 128/129/192/193/256) + extra cached keys for some
    sequences; and the one-token (look-ahead) phase: every sequence has one query row and 1..256 cached keys."""
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    hd = 64
    if phase == "prefill":
        qlens = [1, 2, 17, 64, 127, 128, 129, 190, 256, 33, 100, 192, 193]
        extra = [0, 0, 3, 0, 0, 0, 0, 10, 0, 64, 28, 0, 0]      # keys already in the cache before the first query
    elif phase == "one_token":
        qlens = [1] * 12
        extra = [0, 1, 15, 16, 66, 110, 127, 128, 190, 191, 192, 255]
    else:                                                       # every prompt <= 128 keys: the 4-CTAs-per-SM variant
        qlens = [1] * 10
        extra = [0, 1, 15, 16, 31, 32, 66, 110, 126, 127]
    slack = 10
    q_start, kv_start, t, c = [], [], 0, 0
    for ql, ex in zip(qlens, extra):
        q_start.append(t); kv_start.append(c); t += ql; c += ql + ex + slack
    T, slots = t, c + 300
    qkv_dim = (nh + 2 * nkv) * hd
    qkv = _rand((T, qkv_dim), 1.0, 11)
    kc, vc = _rand((slots, nkv * hd), 1.0, 12), _rand((slots, nkv * hd), 1.0, 13)
    out = torch.zeros((T, nh * hd), dtype=torch.bfloat16, device="cuda")
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device="cuda")
    kvl = [a + b for a, b in zip(qlens, extra)]
    a_qs, a_ql, a_ks, a_kl = i32(q_start), i32(qlens), i32(kv_start), i32(kvl)
    # prompt pass on the tensor cores: 1 = P kept in tensor memory (default), 2 = first kernel (P through shared memory)
    _lib.check(lib.b2s_set_option(b"attn_ts", 0 if use_tc == 2 else 1), "set_option")
    try:
        _lib.check(lib.b2s_attention_bf16(_lib.ptr(qkv), _lib.ptr(kc), _lib.ptr(vc), _lib.ptr(out), _lib.ptr(a_qs), _lib.ptr(a_ql),
                                          _lib.ptr(a_ks), _lib.ptr(a_kl), len(qlens), T, slots, max(qlens), max(kvl), nh, nkv, hd,
                                          qkv_dim, 0, hd, min(use_tc, 1), _lib.current_stream()), "attention")
        torch.cuda.synchronize()
    finally:
        _lib.check(lib.b2s_set_option(b"attn_ts", 1), "set_option")
    q = qkv[:, :nh * hd].view(T, nh, hd)
    ref = _attn_ref(q, kc.view(slots, nkv, hd), vc.view(slots, nkv, hd), list(zip(q_start, qlens, kv_start, kvl)), extra)
    for (qs, ql), r in zip(zip(q_start, qlens), ref):
        got = out[qs:qs + ql].view(ql, nh, hd).float()
        err = (got - r).abs().max().item()
        assert err <= (3e-2 if use_tc else 1.5e-2), (use_tc, ql, err)     # bf16 output (+ bf16 P on the tensor-core path)


@pytest.mark.parametrize("nh,nkv,interleaved,hd", [(4, 2, False, 128), (3, 3, True, 128), (8, 1, False, 128), (6, 6, False, 128),
                                                   (5, 5, True, 80), (4, 4, False, 80), (6, 2, False, 80)])
def test_attention_tc_head_dim_128(nh, nkv, interleaved, hd):
    """Prompt pass on the tensor cores at head_dim 128 (StableLM-alpha / RedPajama-7B / Mistral / BLOOM shapes): K, V and Q
    staged as two 64-wide sub-tiles, grouped-query and per-head [q, k, v] interleaved query layouts, ragged lengths up to
    256 keys with cached keys in front of some prompts.  head_dim 80 (RedPajama-3B) rides the same kernel: a 3D tensor map
    zero-fills columns 80..127 of every Q, K and V box and only 16 columns of the second output slice are stored."""
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    qlens = [1, 2, 17, 64, 127, 128, 129, 190, 256, 33, 100, 192, 193]
    extra = [0, 0, 3, 0, 0, 0, 0, 10, 0, 64, 28, 0, 0]
    q_start, kv_start, t, c = [], [], 0, 0
    for ql, ex in zip(qlens, extra):
        q_start.append(t); kv_start.append(c); t += ql; c += ql + ex + 10
    T, slots = t, c + 300
    if interleaved:                      # HF gpt_neox: per head [q, k, v]
        qkv_dim, q_base, q_hstride = 3 * nh * hd, 0, 3 * hd
    else:
        qkv_dim, q_base, q_hstride = (nh + 2 * nkv) * hd, 0, hd
    qkv = _rand((T, qkv_dim), 1.0, 71)
    kc, vc = _rand((slots, nkv * hd), 1.0, 72), _rand((slots, nkv * hd), 1.0, 73)
    out = torch.zeros((T, nh * hd), dtype=torch.bfloat16, device="cuda")
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device="cuda")
    kvl = [a + b for a, b in zip(qlens, extra)]
    a_qs, a_ql, a_ks, a_kl = i32(q_start), i32(qlens), i32(kv_start), i32(kvl)
    _lib.check(lib.b2s_attention_bf16(_lib.ptr(qkv), _lib.ptr(kc), _lib.ptr(vc), _lib.ptr(out), _lib.ptr(a_qs), _lib.ptr(a_ql),
                                      _lib.ptr(a_ks), _lib.ptr(a_kl), len(qlens), T, slots, max(qlens), max(kvl), nh, nkv, hd,
                                      qkv_dim, q_base, q_hstride, 1, _lib.current_stream()), "attention")
    torch.cuda.synchronize()
    q = qkv.view(T, -1, hd)[:, 0::3][:, :nh] if interleaved else qkv[:, :nh * hd].view(T, nh, hd)
    ref = _attn_ref(q, kc.view(slots, nkv, hd), vc.view(slots, nkv, hd), list(zip(q_start, qlens, kv_start, kvl)), extra)
    for (qs, ql), r in zip(zip(q_start, qlens), ref):
        err = (out[qs:qs + ql].view(ql, nh, hd).float() - r).abs().max().item()
        assert err <= 3e-2, (ql, err)


@pytest.mark.parametrize("name,M,N,K,epi,f16", [("down", 16384, 4544, 18176, 2, True), ("up_gelu", 16384, 18176, 4544, 1, True),
                                                ("down_bf16", 16384, 4544, 18176, 2, False)])
def test_gemm_tc_default_policy_at_bench_sizes(name, M, N, K, epi, f16):
    """The configuration bench.py actually runs, with the LIBRARY DEFAULTS (no gemm_cta_group override): Falcon-7B down
    projection (M >= 4096 and K >= 8192 -> 512x256 CTA-pair tiles, 8 waves -> wave meeting armed) and up projection +
    GELU (256x256 CTA-pair tiles, 60 waves, meeting armed), against fp32 torch on the same 16-bit operands."""
    import gpu_util as U
    dt = torch.float16 if f16 else torch.bfloat16
    A, W = _rand((M, K), 1.0, 31, dt), _rand((N, K), 0.02, 32, dt)
    fn = U.gemm_f16 if f16 else U.gemm_bf16
    if epi == 2:
        resid = _rand((M, N), 1.0, 33, torch.float32)
        out = fn(A, W, None, 2, out=resid.clone())
    else:
        out = fn(A, W, None, epi).float()
    torch.cuda.synchronize()
    worst = 0.0
    for r0 in range(0, M, 2048):                      # reference in row slabs: fp32 copies of A would not fit otherwise
        ref = A[r0:r0 + 2048].float() @ W.float().T
        if epi == 1:
            ref = torch.nn.functional.gelu(ref)
        if epi == 2:
            ref = ref + resid[r0:r0 + 2048]
        worst = max(worst, ((out[r0:r0 + 2048] - ref).abs() / (ref.abs() + 1.0)).max().item())
    tol = 2e-3 if epi == 2 else (2.0 ** -10 if f16 else 2.0 ** -7)
    assert worst <= tol, (name, worst)
    # and bit-identical to the single-CTA tiling without the meeting (the summation order of a row does not depend on either)
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    try:
        _lib.check(lib.b2s_set_option(b"gemm_cta_group", 1), "set_option")
        _lib.check(lib.b2s_set_option(b"gemm_wave_sync", 0), "set_option")
        out1 = fn(A, W, None, 2, out=resid.clone()) if epi == 2 else fn(A, W, None, epi).float()
        torch.cuda.synchronize()
    finally:
        _lib.check(lib.b2s_set_option(b"gemm_cta_group", 0), "set_option")
        _lib.check(lib.b2s_set_option(b"gemm_wave_sync", 4), "set_option")
    assert torch.equal(out, out1), name


def test_gemm_f32_large_matches_fp64():
    """The register-tiled fp32 check GEMM at a Falcon-sized K against an fp64 reference (error of a K = 18176 fmaf chain)."""
    import gpu_util as U
    M, N, K = 300, 520, 18176
    A, W = _rand((M, K), 1.0, 41, torch.float32), _rand((N, K), 0.02, 42, torch.float32)
    out = U.gemm_f32(A, W, None, 4)
    ref = (A.double() @ W.double().T)
    torch.cuda.synchronize()
    err = ((out.double() - ref).abs() / (ref.abs() + 1.0)).max().item()
    assert err <= 5e-5, err      # measured 2.4e-5: what one fp32 fmaf chain of 18176 terms costs


@pytest.mark.parametrize("opt,val", [("gemm_band_n", 3), ("gemm_band_n", 0), ("gemm_prefetch", 8), ("gemm_prefetch", 3)])
@pytest.mark.parametrize("M,N,K,epi", [(9000, 4544, 4544, 2), (5000, 2304, 832, 1)])
def test_gemm_tc_rasterisation_and_prefetch_knobs_keep_the_bits(opt, val, M, N, K, epi):
    """Column-band tile order and the producers' L2 prefetch change when a tile is computed and what is in L2, never a
    result: the output must equal the default configuration's bit for bit (several bands, several waves, ragged edges)."""
    import gpu_util as U
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    A, W = _rand((M, K), 1.0, 51, torch.float16), _rand((N, K), 0.03, 52, torch.float16)
    resid = _rand((M, N), 1.0, 53, torch.float32)
    run = lambda: U.gemm_f16(A, W, None, 2, out=resid.clone()) if epi == 2 else U.gemm_f16(A, W, None, epi)
    base = run()
    torch.cuda.synchronize()
    try:
        _lib.check(lib.b2s_set_option(opt.encode(), val), "set_option")
        got = run()
        torch.cuda.synchronize()
    finally:
        _lib.check(lib.b2s_set_option(opt.encode(), -1 if opt == "gemm_band_n" else 0), "set_option")     # library defaults
    assert torch.equal(base, got), (opt, val)
    ref = A.float() @ W.float().T
    ref = ref + resid if epi == 2 else torch.nn.functional.gelu(ref)
    assert ((got.float() - ref).abs() / (ref.abs() + 1.0)).max().item() <= 2e-3


@pytest.mark.parametrize("M,ff,K", [(300, 704, 256), (1000, 14336, 512), (77, 1152, 520)])
@pytest.mark.parametrize("f16", [False, True])
def test_gemm_tc_swiglu_epilogue(M, ff, K, f16, cta_group):
    """SwiGLU formed in the epilogue: W rows alternate gate_i, up_i; out[M, ff] = silu(gate) * up (HF llama LlamaMLP)."""
    import gpu_util as U
    dt = torch.float16 if f16 else torch.bfloat16
    A = _rand((M, K), 1.0, 61, dt)
    Wg, Wu = _rand((ff, K), 0.05, 62, dt), _rand((ff, K), 0.05, 63, dt)
    W = torch.stack((Wg, Wu), dim=1).reshape(2 * ff, K).contiguous()
    out = torch.empty((M, ff), dtype=dt, device="cuda")
    (U.gemm_f16 if f16 else U.gemm_bf16)(A, W, None, 6, out=out)
    torch.cuda.synchronize()
    ref = torch.nn.functional.silu(A.float() @ Wg.float().T) * (A.float() @ Wu.float().T)
    err = ((out.float() - ref).abs() / (ref.abs() + 1.0)).max().item()
    assert err <= (2.0 ** -10 if f16 else 2.0 ** -7), err
