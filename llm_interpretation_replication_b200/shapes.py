"""Decoder shapes for the scoring path.

The reference never names shapes itself: it calls ``AutoModelForCausalLM.from_pretrained``
(reference ``analysis/run_base_vs_instruct_100q.py:426-450``).  The shapes below are the
ones SURVEY.md Appendix C records (Falcon-7B = ``transformers`` ``FalconConfig()`` defaults;
the GPT-NeoX ones recalled from the public ``config.json`` files) plus tiny variants of each
family that the parity tests run against the HF CPU path.
"""
from __future__ import annotations

from dataclasses import dataclass, asdict, replace

# arch codes (must match include/b200score.h)
ARCH_FALCON = 0
ARCH_GPT_NEOX = 1
ARCH_LLAMA = 2            # Llama / Mistral family: RMSNorm, SwiGLU, grouped-query attention, no biases
ARCH_OPT = 3              # OPT: learned positions (offset 2), ReLU MLP, pre-LayerNorm, biases, no rotary
ARCH_BLOOM = 4            # BLOOM: ALiBi, LayerNorm on the embeddings, tanh-GELU MLP, per-head [q,k,v] fused projection

# position encodings and MLP activations (must match include/b200score.h)
POS_ROPE = 0
POS_LEARNED = 1           # x = wte[token] + wpe[position + pos_offset]          (HF opt OPTLearnedPositionalEmbedding)
POS_ALIBI = 2             # score(q, k) += slope[head] * key_position             (HF bloom build_alibi_tensor)
ACT_GELU = 0              # exact (erf) GELU
ACT_RELU = 1
ACT_GELU_TANH = 2         # 0.5 x (1 + tanh(0.79788456 x (1 + 0.044715 x^2)))     (HF bloom bloom_gelu_forward)

# normalisation and MLP kinds (must match include/b200score.h)
NORM_LAYERNORM = 0
NORM_RMS = 1              # x * rsqrt(mean(x^2) + eps) * g, no bias   (HF llama LlamaRMSNorm)
MLP_GELU = 0              # down(gelu(up(x)))
MLP_SWIGLU = 1            # down(silu(gate(x)) * up(x)); the W_UP slot holds [gate rows; up rows]  (HF llama LlamaMLP)
MLP_SWIGLU_FUSED = 2      # the same MLP with the W_UP rows interleaved (gate_0, up_0, gate_1, ...): what Scorer hands to the library in
                          # the tensor-core modes, so that silu(gate) * up is formed in the GEMM epilogue

# qkv column layouts (must match include/b200score.h)
QKV_FALCON_MQA = 0       # [q_0 .. q_{nh-1}, k_0 .. k_{nkv-1}, v_0 .. v_{nkv-1}] each head_dim wide (HF falcon _split_heads with
                         # multi_query: nkv = 1; also cat(q_proj, k_proj, v_proj) of the Llama family)
QKV_NEOX_INTERLEAVED = 1  # per head [q, k, v]                            (HF gpt_neox qkv.chunk(3))

# compute modes (must match include/b200score.h)
MODE_BF16_FAST = 0    # bf16 MMA operands, fp32 accumulate in TMEM, fp32 residual stream
MODE_FP16 = 1         # fp16 weights + activations on the tensor cores (holds the 2e-3 gate at full depth)
MODE_BF16_SPLIT = 2   # bf16 weights; GEMM activations as hi + lo bf16 halves (holds the gate with bf16 range, 2x tensor work)
MODE_FP32_CHECK = 3   # fp32 weights + activations on CUDA cores; parity gate 1e-5
MODE_FP16_FFN_SPLIT = 5   # T5 handles only: fp16 everywhere, the feed-forward activation as bf16 hi + lo halves against bf16 wo
MODE_BF16_SPLIT_ATTN32 = 6   # decoder-only handles: BF16_SPLIT with q/k/v, the KV cache and attention in fp32 (CUDA-core kernels)
MODE_INT8_EMU = 4     # bitsandbytes LLM.int8 (the reference's USE_8BIT configuration) emulated on the FP16 kernels


@dataclass(frozen=True)
class DecoderShape:
    name: str
    arch: int
    n_layer: int
    d_model: int
    n_head: int
    n_kv_head: int
    head_dim: int
    d_ff: int
    vocab: int
    rotary_dim: int
    rope_theta: float = 10000.0
    norm_eps: float = 1e-5
    parallel_residual: bool = True
    n_ln_per_layer: int = 1
    has_bias: bool = False
    tied_head: bool = True
    qkv_layout: int = QKV_FALCON_MQA
    eos_id: int = 11
    max_pos: int = 2048
    norm_type: int = NORM_LAYERNORM
    mlp_type: int = MLP_GELU
    pos_type: int = POS_ROPE
    pos_offset: int = 0          # rows of wpe skipped before position 0 (OPT: 2)
    act_type: int = ACT_GELU
    emb_norm: bool = False       # LayerNorm applied to the token embeddings (BLOOM word_embeddings_layernorm)
    qkv_bias: bool = False       # a bias on the fused q|k|v projection only (Qwen: c_attn / q,k,v_proj biases; has_bias stays False)

    @property
    def up_dim(self) -> int:
        """Rows of the W_UP slot: d_ff, or 2 * d_ff (gate rows, then up rows) for SwiGLU."""
        return self.d_ff * (2 if self.mlp_type == MLP_SWIGLU else 1)

    @property
    def qkv_dim(self) -> int:
        return (self.n_head + 2 * self.n_kv_head) * self.head_dim

    @property
    def layer_params(self) -> int:
        d, ff = self.d_model, self.d_ff
        return d * self.qkv_dim + d * self.n_head * self.head_dim + d * self.up_dim + d * ff

    def flops_per_token(self) -> float:
        """Algorithmic GEMM FLOPs per prompt token."""
        return 2.0 * self.n_layer * self.layer_params

    def flops_per_position(self) -> float:
        """LM-head FLOPs per scored position."""
        return 2.0 * self.vocab * self.d_model

    def attn_flops(self, L: int) -> float:
        """Causal attention FLOPs for one prompt of L tokens (QK^T and PV, half masked)."""
        return 4.0 * L * L * self.n_head * self.head_dim * self.n_layer / 2.0

    def to_dict(self):
        return asdict(self)

    def with_layers(self, n):
        return replace(self, n_layer=n, name=f"{self.name}-L{n}")


def falcon_7b() -> DecoderShape:
    # HF:models/falcon/configuration_falcon.py:53-76
    return DecoderShape("falcon-7b", ARCH_FALCON, 32, 4544, 71, 1, 64, 18176, 65024, 64,
                        parallel_residual=True, n_ln_per_layer=1, has_bias=False, tied_head=True,
                        qkv_layout=QKV_FALCON_MQA, eos_id=11)


def stablelm_alpha_7b() -> DecoderShape:
    return DecoderShape("stablelm-base-alpha-7b", ARCH_GPT_NEOX, 16, 6144, 48, 48, 128, 24576, 50432, 32,
                        parallel_residual=True, n_ln_per_layer=2, has_bias=True, tied_head=False,
                        qkv_layout=QKV_NEOX_INTERLEAVED, eos_id=0, max_pos=4096)


def redpajama_7b() -> DecoderShape:
    return DecoderShape("redpajama-incite-7b", ARCH_GPT_NEOX, 32, 4096, 32, 32, 128, 16384, 50432, 128,
                        parallel_residual=False, n_ln_per_layer=2, has_bias=True, tied_head=False,
                        qkv_layout=QKV_NEOX_INTERLEAVED, eos_id=0)


def redpajama_3b() -> DecoderShape:
    return DecoderShape("redpajama-incite-base-3b", ARCH_GPT_NEOX, 32, 2560, 32, 32, 80, 10240, 50432, 80,
                        parallel_residual=False, n_ln_per_layer=2, has_bias=True, tied_head=False,
                        qkv_layout=QKV_NEOX_INTERLEAVED, eos_id=0)


def h2ogpt_12b() -> DecoderShape:
    return DecoderShape("h2ogpt-oasst1-512-12b", ARCH_GPT_NEOX, 36, 5120, 40, 40, 128, 20480, 50688, 32,
                        parallel_residual=True, n_ln_per_layer=2, has_bias=True, tied_head=False,
                        qkv_layout=QKV_NEOX_INTERLEAVED, eos_id=0)


def dolly_v2_7b() -> DecoderShape:
    # pythia-6.9b shape
    return DecoderShape("dolly-v2-7b", ARCH_GPT_NEOX, 32, 4096, 32, 32, 128, 16384, 50280, 32,
                        parallel_residual=True, n_ln_per_layer=2, has_bias=True, tied_head=False,
                        qkv_layout=QKV_NEOX_INTERLEAVED, eos_id=0)


def mistral_7b() -> DecoderShape:
    # mistralai/Mistral-7B-Instruct-v0.3 (compare_instruct_models.py:157): GQA 32/8, head_dim 128, SwiGLU 14336,
    # vocab 32768, rope_theta 1e6, no sliding window
    return DecoderShape("mistral-7b", ARCH_LLAMA, 32, 4096, 32, 8, 128, 14336, 32768, 128, rope_theta=1000000.0,
                        norm_eps=1e-5, parallel_residual=False, n_ln_per_layer=2, has_bias=False, tied_head=False,
                        qkv_layout=QKV_FALCON_MQA, eos_id=2, max_pos=32768, norm_type=NORM_RMS, mlp_type=MLP_SWIGLU)


def mistral_7b_v01() -> DecoderShape:
    # mistralai/Mistral-7B-v0.1 (run_base_vs_instruct_100q.py:102, compare_base_vs_instruct.py:155): vocab 32000, rope_theta 1e4.
    # Its 4096-token sliding window never binds on this path (prompts + 50 generated tokens stay far below it).
    return replace(mistral_7b(), name="mistral-7b-v0.1", vocab=32000, rope_theta=10000.0)


def mistral_7b_v02() -> DecoderShape:
    # mistralai/Mistral-7B-Instruct-v0.2: vocab 32000, rope_theta 1e6, no sliding window
    return replace(mistral_7b(), name="mistral-7b-instruct-v0.2", vocab=32000)


def qwen_7b() -> DecoderShape:
    # Qwen/Qwen-7B and Qwen/Qwen-7B-Chat (compare_instruct_models.py:159, compare_base_vs_instruct.py:159): the first-generation
    # Qwen, shipped as custom modelling code (QWenLMHeadModel, trust_remote_code).  Its published forward is the Llama block
    # with a bias on the fused c_attn projection: RMSNorm 1e-6, rotate_half RoPE over the whole 128-wide head (base 10000; the
    # dynamic-NTK and log-n scalings only act beyond seq_length, far above this path's prompts), c_proj(w1(x) * silu(w2(x)))
    # with w1 / w2 of intermediate_size / 2 = 11008 rows, untied lm_head, vocab 151936.
    return DecoderShape("qwen-7b", ARCH_LLAMA, 32, 4096, 32, 32, 128, 11008, 151936, 128, rope_theta=10000.0,
                        norm_eps=1e-6, parallel_residual=False, n_ln_per_layer=2, has_bias=False, tied_head=False,
                        qkv_layout=QKV_FALCON_MQA, eos_id=151643, max_pos=8192, norm_type=NORM_RMS, mlp_type=MLP_SWIGLU,
                        qkv_bias=True)


def baichuan2_7b() -> DecoderShape:
    # baichuan-inc/Baichuan2-7B-Chat (compare_instruct_models.py:147): custom modelling code (BaichuanForCausalLM); the published
    # forward is the Llama block with the q, k, v projections packed in W_pack and an lm_head whose rows are L2-normalised at
    # inference (NormHead) -- done once at load time here.  vocab 125696, RMSNorm 1e-6.
    return DecoderShape("baichuan2-7b", ARCH_LLAMA, 32, 4096, 32, 32, 128, 11008, 125696, 128, rope_theta=10000.0,
                        norm_eps=1e-6, parallel_residual=False, n_ln_per_layer=2, has_bias=False, tied_head=False,
                        qkv_layout=QKV_FALCON_MQA, eos_id=2, max_pos=4096, norm_type=NORM_RMS, mlp_type=MLP_SWIGLU)


def llama2_7b() -> DecoderShape:
    # meta-llama/Llama-2-7b-chat-hf (commented out in the reference's list, compare_instruct_models.py:155)
    return DecoderShape("llama-2-7b", ARCH_LLAMA, 32, 4096, 32, 32, 128, 11008, 32000, 128, rope_theta=10000.0,
                        norm_eps=1e-5, parallel_residual=False, n_ln_per_layer=2, has_bias=False, tied_head=False,
                        qkv_layout=QKV_FALCON_MQA, eos_id=2, max_pos=4096, norm_type=NORM_RMS, mlp_type=MLP_SWIGLU)


def opt_iml_1_3b() -> DecoderShape:
    # facebook/opt-iml-1.3b (compare_instruct_models.py:151): HF:models/opt/configuration_opt.py, opt-1.3b sizes
    return DecoderShape("opt-iml-1.3b", ARCH_OPT, 24, 2048, 32, 32, 64, 8192, 50272, 0, norm_eps=1e-5, parallel_residual=False,
                        n_ln_per_layer=2, has_bias=True, tied_head=True, qkv_layout=QKV_FALCON_MQA, eos_id=2, max_pos=2048,
                        pos_type=POS_LEARNED, pos_offset=2, act_type=ACT_RELU)


def bloomz_7b1() -> DecoderShape:
    # bigscience/bloomz-7b1 (compare_instruct_models.py:148): HF:models/bloom/configuration_bloom.py, bloom-7b1 sizes
    return DecoderShape("bloomz-7b1", ARCH_BLOOM, 30, 4096, 32, 32, 128, 16384, 250880, 0, norm_eps=1e-5, parallel_residual=False,
                        n_ln_per_layer=2, has_bias=True, tied_head=True, qkv_layout=QKV_NEOX_INTERLEAVED, eos_id=2, max_pos=2048,
                        pos_type=POS_ALIBI, act_type=ACT_GELU_TANH, emb_norm=True)


# tiny variants (CPU-oracle sized) --------------------------------------------------------
def tiny_opt(vocab=272, n_layer=2, n_head=3) -> DecoderShape:
    hd = 64
    d = n_head * hd
    return DecoderShape("tiny-opt", ARCH_OPT, n_layer, d, n_head, n_head, hd, 4 * d, vocab, 0, norm_eps=1e-5,
                        parallel_residual=False, n_ln_per_layer=2, has_bias=True, tied_head=True, qkv_layout=QKV_FALCON_MQA,
                        eos_id=2, max_pos=512, pos_type=POS_LEARNED, pos_offset=2, act_type=ACT_RELU)


def tiny_bloom(vocab=280, n_layer=2, n_head=6) -> DecoderShape:
    # 6 heads: not a power of two, so the ALiBi slopes use both branches of HF's formula
    hd = 64
    d = n_head * hd
    return DecoderShape("tiny-bloom", ARCH_BLOOM, n_layer, d, n_head, n_head, hd, 4 * d, vocab, 0, norm_eps=1e-5,
                        parallel_residual=False, n_ln_per_layer=2, has_bias=True, tied_head=True,
                        qkv_layout=QKV_NEOX_INTERLEAVED, eos_id=2, max_pos=2048, pos_type=POS_ALIBI, act_type=ACT_GELU_TANH,
                        emb_norm=True)


def tiny_llama(vocab=288, n_layer=2, n_head=4, n_kv_head=2) -> DecoderShape:
    # head_dim 64: the tensor-core attention and the fused QKV+RoPE epilogue with grouped-query heads
    hd = 64
    d = n_head * hd
    return DecoderShape("tiny-llama", ARCH_LLAMA, n_layer, d, n_head, n_kv_head, hd, 3 * d - 64, vocab, hd, rope_theta=10000.0,
                        norm_eps=1e-5, parallel_residual=False, n_ln_per_layer=2, has_bias=False, tied_head=False,
                        qkv_layout=QKV_FALCON_MQA, eos_id=2, norm_type=NORM_RMS, mlp_type=MLP_SWIGLU)


def tiny_qwen(vocab=304, n_layer=2, n_head=3) -> DecoderShape:
    # Llama block with a q|k|v bias (Qwen): head_dim 64, multi-head
    hd = 64
    d = n_head * hd
    return DecoderShape("tiny-qwen", ARCH_LLAMA, n_layer, d, n_head, n_head, hd, 2 * d + 64, vocab, hd, rope_theta=10000.0,
                        norm_eps=1e-6, parallel_residual=False, n_ln_per_layer=2, has_bias=False, tied_head=False,
                        qkv_layout=QKV_FALCON_MQA, eos_id=2, norm_type=NORM_RMS, mlp_type=MLP_SWIGLU, qkv_bias=True)


def tiny_mistral(vocab=264, n_layer=2, n_head=4, n_kv_head=1) -> DecoderShape:
    # Mistral-like: head_dim 128 (CUDA-core attention, separate RoPE kernel), rope_theta 1e6
    hd = 128
    d = n_head * hd
    return DecoderShape("tiny-mistral", ARCH_LLAMA, n_layer, d, n_head, n_kv_head, hd, 2 * d + 128, vocab, hd,
                        rope_theta=1000000.0, norm_eps=1e-5, parallel_residual=False, n_ln_per_layer=2, has_bias=False,
                        tied_head=False, qkv_layout=QKV_FALCON_MQA, eos_id=2, norm_type=NORM_RMS, mlp_type=MLP_SWIGLU)



def tiny_falcon(vocab=256, n_layer=2, n_head=3) -> DecoderShape:
    hd = 64
    d = n_head * hd
    return DecoderShape("tiny-falcon", ARCH_FALCON, n_layer, d, n_head, 1, hd, 4 * d, vocab, hd,
                        parallel_residual=True, n_ln_per_layer=1, has_bias=False, tied_head=True,
                        qkv_layout=QKV_FALCON_MQA, eos_id=11)


def tiny_neox_parallel(vocab=320, n_layer=2, n_head=2) -> DecoderShape:
    # StableLM-alpha-like: hd 128, rotary_pct 0.25, parallel residual, biases
    hd = 128
    d = n_head * hd
    return DecoderShape("tiny-neox-par", ARCH_GPT_NEOX, n_layer, d, n_head, n_head, hd, 4 * d, vocab, 32,
                        parallel_residual=True, n_ln_per_layer=2, has_bias=True, tied_head=False,
                        qkv_layout=QKV_NEOX_INTERLEAVED, eos_id=0)


def tiny_neox_sequential(vocab=200, n_layer=2, n_head=2) -> DecoderShape:
    # RedPajama-3B-like: hd 80, full rotary, sequential residual, biases
    hd = 80
    d = n_head * hd
    return DecoderShape("tiny-neox-seq", ARCH_GPT_NEOX, n_layer, d, n_head, n_head, hd, 4 * d, vocab, hd,
                        parallel_residual=False, n_ln_per_layer=2, has_bias=True, tied_head=False,
                        qkv_layout=QKV_NEOX_INTERLEAVED, eos_id=0)


REGISTRY = {
    "falcon-7b": falcon_7b,
    "stablelm-base-alpha-7b": stablelm_alpha_7b,
    "redpajama-incite-7b": redpajama_7b,
    "redpajama-incite-base-3b": redpajama_3b,
    "h2ogpt-oasst1-512-12b": h2ogpt_12b,
    "dolly-v2-7b": dolly_v2_7b,
    "mistral-7b": mistral_7b,
    "mistral-7b-v0.1": mistral_7b_v01,
    "mistral-7b-instruct-v0.2": mistral_7b_v02,
    "llama-2-7b": llama2_7b,
    "qwen-7b": qwen_7b,
    "baichuan2-7b": baichuan2_7b,
    "tiny-qwen": tiny_qwen,
    "opt-iml-1.3b": opt_iml_1_3b,
    "bloomz-7b1": bloomz_7b1,
    "tiny-opt": tiny_opt,
    "tiny-bloom": tiny_bloom,
    "tiny-llama": tiny_llama,
    "tiny-mistral": tiny_mistral,
    "tiny-falcon": tiny_falcon,
    "tiny-neox-par": tiny_neox_parallel,
    "tiny-neox-seq": tiny_neox_sequential,
}

# HF hub name -> shape, for the kept entry points (random-init; no weights offline)
HUB_TO_SHAPE = {
    "tiiuae/falcon-7b": falcon_7b, "tiiuae/falcon-7b-instruct": falcon_7b,
    "stabilityai/stablelm-base-alpha-7b": stablelm_alpha_7b,
    "stabilityai/stablelm-tuned-alpha-7b": stablelm_alpha_7b,
    "togethercomputer/RedPajama-INCITE-7B-Base": redpajama_7b,
    "togethercomputer/RedPajama-INCITE-7B-Instruct": redpajama_7b,
    "togethercomputer/RedPajama-INCITE-Base-3B-v1": redpajama_3b,
    "togethercomputer/RedPajama-INCITE-Instruct-3B-v1": redpajama_3b,
    "h2oai/h2ogpt-oasst1-512-12b": h2ogpt_12b,
    "EleutherAI/pythia-6.9b": dolly_v2_7b, "databricks/dolly-v2-7b": dolly_v2_7b,
    "facebook/opt-iml-1.3b": opt_iml_1_3b, "facebook/opt-1.3b": opt_iml_1_3b,
    "bigscience/bloomz-7b1": bloomz_7b1, "bigscience/bloom-7b1": bloomz_7b1,
    "mistralai/Mistral-7B-Instruct-v0.3": mistral_7b, "mistralai/Mistral-7B-v0.3": mistral_7b,
    "mistralai/Mistral-7B-v0.1": mistral_7b_v01, "mistralai/Mistral-7B-Instruct-v0.2": mistral_7b_v02,
    "meta-llama/Llama-2-7b-chat-hf": llama2_7b, "meta-llama/Llama-2-7b-hf": llama2_7b,
    "Qwen/Qwen-7B": qwen_7b, "Qwen/Qwen-7B-Chat": qwen_7b, "baichuan-inc/Baichuan2-7B-Chat": baichuan2_7b,
}
