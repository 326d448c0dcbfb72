"""The allocation-free LM-head entry point (b2s_lm_head_bf16_static) against the regular one.

Some users of the library only ever get input and output pointers -- a plug-in inside an application whose
memory pool already owns the GPU, or a serving loop that must never hit the device-wide stall of
cudaMalloc -- so this variant keeps its scratch in a fixed array inside the library.  The test calls it twice
in a row with different inputs (the second call finds the first call's leftovers in the shared array) and
expects the same statistics as the entry point that takes caller-provided scratch.  It lives in its own file,
collected last, and must pass both on its own and after every other test has used the library.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-3   # the tolerance test_lm_head_fused uses for the same statistics


def _case(M, V, d, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    h = torch.randn((M, d), generator=g, device="cuda").to(torch.bfloat16)
    W = (torch.randn((V, d), generator=g, device="cuda") * 0.05).to(torch.bfloat16)
    cand = torch.randint(0, V, (M, 2), generator=g, device="cuda", dtype=torch.int32)
    return h, W, cand


def _static(lib, h, W, cand):
    from llm_interpretation_replication_b200 import _lib
    M, d = h.shape
    stats = torch.full((M, 16), float("nan"), dtype=torch.float32, device="cuda")
    _lib.check(lib.b2s_lm_head_bf16_static(_lib.ptr(h), _lib.ptr(W), _lib.ptr(cand), M, W.shape[0], d, _lib.ptr(stats),
                                           _lib.current_stream()), "b2s_lm_head_bf16_static")
    return stats


def test_static_scratch_lm_head_matches_regular_entry_point_twice():
    import gpu_util as U
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    assert lib.b2s_lm_head_static_capacity() == (4 << 20) + 4096
    # two different problems back to back: the second one is smaller, so stale partials of the first remain
    # in the array beyond what it rewrites
    for (M, V, d, seed) in [(200, 50432, 512, 1), (37, 1000, 192, 2)]:
        h, W, cand = _case(M, V, d, seed)
        got = _static(lib, h, W, cand)
        want = U.lm_head(h, W, cand)
        torch.cuda.synchronize()
        assert torch.isfinite(got[:, :9]).all()
        assert (got[:, :9] - want[:, :9]).abs().max().item() <= TOL
        assert torch.equal(got.view(torch.int32)[:, 15], want.view(torch.int32)[:, 15])   # arg max index
    # same inputs twice in a row give identical results (nothing left over from the previous call leaks in)
    h, W, cand = _case(64, 65024, 256, 3)
    a = _static(lib, h, W, cand)
    b = _static(lib, h, W, cand)
    torch.cuda.synchronize()
    assert torch.equal(a, b)


def test_static_scratch_rejects_oversized_problem():
    from llm_interpretation_replication_b200 import _lib
    lib = _lib.load()
    M, V, d = 1024, 65024, 64            # needs ~8.3 MB of partials: over the built-in 4 MiB
    assert lib.b2s_lm_head_scratch_bytes(M, V) > lib.b2s_lm_head_static_capacity()
    h, W, cand = _case(M, 1024, d, 4)    # small real buffers: the call must refuse before touching them
    stats = torch.zeros((M, 16), dtype=torch.float32, device="cuda")
    rc = lib.b2s_lm_head_bf16_static(_lib.ptr(h), _lib.ptr(W), _lib.ptr(cand), M, V, d, _lib.ptr(stats), _lib.current_stream())
    assert rc != 0
    msg = lib.b2s_last_error().decode()
    assert "scratch" in msg and "holds" in msg
    torch.cuda.synchronize()
    assert float(stats.abs().sum()) == 0.0
