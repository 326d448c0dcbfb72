"""LLM.int8 emulation on CPU: the restated Linear8bitLt (oracle/int8_ref.py) against hand-computed cases and against the
fake-quantised formulation the CUDA path uses (operands put on the int8 grid, multiplied as fp16 x fp16 -> fp32)."""
import torch

from oracle import int8_ref, decoder_ref
from llm_interpretation_replication_b200 import shapes, weights
from llm_interpretation_replication_b200.scorer import int8_rowwise_dequant


def test_linear8bit_hand_case():
    a = torch.tensor([[1.0, -2.0, 0.5, 4.0], [8.0, 1.0, -1.0, 0.25]])          # 8.0 >= 6: column 0 is an outlier column
    w = torch.tensor([[0.5, -1.0, 0.25, 1.0], [2.0, 0.0, -0.5, 0.1]])
    out = int8_ref.linear8bit(a, w, out_fp16=False)
    cb, scb = int8_ref.quantize_weight(w)
    assert scb.tolist() == [1.0, 2.0] and cb[0].tolist() == [64.0, -127.0, 32.0, 127.0]
    # row 0: statistics over all four elements (none >= 6): SCA = 4; row 1: 8.0 is left out: SCA = 1
    ca0 = torch.round(a[0] * 127 / 4.0); ca0[0] = 0
    ca1 = torch.round(a[1] * 127 / 1.0); ca1[0] = 0
    want = torch.stack([ca0 @ cb.t() * 4.0, ca1 @ cb.t() * 1.0]) * scb[None, :] / (127.0 * 127.0)
    want = want + a[:, :1] @ (cb[:, :1] * scb[:, None] / 127.0).half().float().t()
    assert torch.allclose(out, want, atol=1e-6)
    # threshold 0: no decomposition, every element on the grid
    out0 = int8_ref.linear8bit(a, w, threshold=0.0, out_fp16=False)
    ca = torch.round(a * 127 / a.abs().amax(1, keepdim=True))
    assert torch.allclose(out0, (ca @ cb.t()) * a.abs().amax(1, keepdim=True) * scb[None, :] / (127.0 * 127.0), atol=1e-6)


def test_fake_quant_formulation_is_within_fp16_rounding_of_int8():
    """What the CUDA path multiplies (dequantised operands in fp16) against the exact integer product: the difference is
    the fp16 rounding of the operands, an order of magnitude below the int8 quantisation error itself."""
    g = torch.Generator().manual_seed(0)
    a = (torch.randn(33, 256, generator=g) * 1.5).half().float()
    a[3, 17] = 9.5; a[20, 200] = -6.25
    w = torch.randn(96, 256, generator=g) * 0.04
    ref = int8_ref.linear8bit(a, w, out_fp16=False)
    big = a.abs() >= 6.0
    idx = big.any(0)
    sca = torch.where(big, torch.zeros_like(a), a.abs()).amax(1, keepdim=True)
    aq = (torch.round(a * (127 / sca)) * (sca / 127)).half().float()
    aq[:, idx] = a[:, idx]
    emu = aq @ int8_rowwise_dequant(w).float().t()
    exact = a @ w.t()
    e_emu, e_q = (ref - emu).abs().max().item(), (ref - exact).abs().max().item()
    assert e_emu < 2e-3 and e_q > 10 * e_emu, (e_emu, e_q)


def test_call_structure_matches_cached_generation():
    """Rows cut as [prompt | one row per generated token] reproduce, position by position, what a cached decode computes:
    the logits of position t do not change when later rows are appended."""
    shp = shapes.tiny_falcon()
    W = weights.init_weights(shp, seed=5)
    f16 = lambda t: t.to(torch.float16).to(t.dtype)
    ids = torch.randint(32, shp.vocab - 32, (21,), generator=torch.Generator().manual_seed(2))
    fn = int8_ref.make_linear_fn(f16, prompt_len=21)
    h_a = decoder_ref.forward_hidden(shp, W, ids, round_act=f16, linear_fn=fn)
    ids2 = torch.cat([ids, torch.tensor([40, 41, 42])])
    h_b = decoder_ref.forward_hidden(shp, W, ids2, round_act=f16, linear_fn=fn)
    assert torch.equal(h_a, h_b[:21])
    fn_one = int8_ref.make_linear_fn(f16, prompt_len=24)       # the same rows as ONE call: other outlier columns / a different result
    h_c = decoder_ref.forward_hidden(shp, W, ids2, round_act=f16, linear_fn=fn_one)
    assert h_c.shape == h_b.shape


def test_llm_int8_disagrees_with_itself_at_its_own_noise_level():
    """Why the GPU emulation is held to a statistical bound: run the restated LLM.int8 twice, the second time with every
    activation that enters a linear layer moved by at most one fp16 ulp (what another correct implementation of the layers in
    between would hand over).  rint() turns such a difference into a whole int8 step for the elements near a rounding boundary,
    and relative_prob moves by as much as int8 moves it away from fp16 in the first place; the fp16 model under the same
    jitter barely moves."""
    shp = shapes.tiny_neox_sequential()
    W = weights.init_weights(shp, seed=103)
    f16 = lambda t: t.to(torch.float16).to(t.dtype)
    gen = torch.Generator().manual_seed(0)

    def jitter(t):
        h = t.to(torch.float16)
        step = (torch.rand(h.shape, generator=gen) < 0.5).to(torch.int16) * torch.where(torch.rand(h.shape, generator=gen) < 0.5, 1, -1).to(torch.int16)
        return (h.view(torch.int16) + step).view(torch.float16).float().to(t.dtype)

    g = torch.Generator().manual_seed(4)
    d_self, d_q, d_f16 = [], [], []
    for L in (40, 75, 110, 64, 90, 33):
        ids = torch.randint(32, shp.vocab - 32, (L,), generator=g)
        lo8 = decoder_ref.last_logits(shp, W, ids, round_act=f16, linear_fn=int8_ref.make_linear_fn(f16, L))
        lo8j = decoder_ref.last_logits(shp, W, ids, round_act=jitter, linear_fn=int8_ref.make_linear_fn(jitter, L))
        lo16 = decoder_ref.last_logits(shp, W, ids, round_act=f16)
        lo16j = decoder_ref.last_logits(shp, W, ids, round_act=jitter)
        top = torch.topk(lo16, 2).indices

        def rel(lo):
            p = torch.softmax(lo, -1)
            return (p[top[0]] / (p[top[0]] + p[top[1]])).item()
        d_self.append(abs(rel(lo8) - rel(lo8j))); d_q.append(abs(rel(lo8) - rel(lo16))); d_f16.append(abs(rel(lo16) - rel(lo16j)))
    m_self, m_q, m_f16 = (sum(v) / len(v) for v in (d_self, d_q, d_f16))
    assert m_self > 0.3 * m_q and m_f16 < 0.3 * m_q, (m_self, m_q, m_f16)
