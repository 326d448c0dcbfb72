"""HF ``transformers`` CPU models over the flat weight layout (test infrastructure only; see oracle/__init__.py).

``build_hf_model`` binds the same tensors the C-ABI library receives (``llm_interpretation_replication_b200.weights``)
to the HF model class the reference would load with ``AutoModelForCausalLM.from_pretrained``
(analysis/run_base_vs_instruct_100q.py:426-450), so the oracle and the CUDA path see identical numbers.  Used by the
golden generator, the parity tests, ``tools/`` and the CPU legs of ``bench.py`` -- never by the product package.
"""
from __future__ import annotations

from typing import Dict

import torch

from llm_interpretation_replication_b200.shapes import (DecoderShape, ARCH_FALCON, ARCH_GPT_NEOX, ARCH_LLAMA, ARCH_OPT,
                                                        ARCH_BLOOM)


def hf_config(shape: DecoderShape):
    """HF config equivalent to ``shape`` (used by the oracle and the golden generator only)."""
    if shape.arch == ARCH_FALCON:
        from transformers import FalconConfig
        return FalconConfig(vocab_size=shape.vocab, hidden_size=shape.d_model,
                            num_hidden_layers=shape.n_layer, num_attention_heads=shape.n_head,
                            multi_query=True, parallel_attn=True, bias=False, alibi=False,
                            new_decoder_architecture=False, layer_norm_epsilon=shape.norm_eps,
                            bos_token_id=shape.eos_id, eos_token_id=shape.eos_id,
                            max_position_embeddings=shape.max_pos, tie_word_embeddings=True,
                            hidden_dropout=0.0, attention_dropout=0.0)
    if shape.arch == ARCH_GPT_NEOX:
        from transformers import GPTNeoXConfig
        return GPTNeoXConfig(vocab_size=shape.vocab, hidden_size=shape.d_model,
                             num_hidden_layers=shape.n_layer, num_attention_heads=shape.n_head,
                             intermediate_size=shape.d_ff, hidden_act="gelu",
                             rotary_pct=shape.rotary_dim / shape.head_dim, rotary_emb_base=shape.rope_theta,
                             layer_norm_eps=shape.norm_eps, use_parallel_residual=shape.parallel_residual,
                             bos_token_id=shape.eos_id, eos_token_id=shape.eos_id,
                             max_position_embeddings=shape.max_pos, tie_word_embeddings=False,
                             attention_bias=True, hidden_dropout=0.0, attention_dropout=0.0)
    if shape.arch == ARCH_LLAMA and shape.qkv_bias:
        # The Llama block with q/k/v biases is HF's Qwen2 (HF:models/qwen2/modeling_qwen2.py: q_proj/k_proj/v_proj bias=True,
        # o_proj and the MLP without).  The first-generation Qwen of the reference's lists is custom code with the same math.
        from transformers import Qwen2Config
        return Qwen2Config(vocab_size=shape.vocab, hidden_size=shape.d_model, intermediate_size=shape.d_ff,
                           num_hidden_layers=shape.n_layer, num_attention_heads=shape.n_head,
                           num_key_value_heads=shape.n_kv_head, hidden_act="silu", max_position_embeddings=shape.max_pos,
                           rms_norm_eps=shape.norm_eps, rope_theta=shape.rope_theta, use_sliding_window=False,
                           tie_word_embeddings=shape.tied_head, bos_token_id=1, eos_token_id=shape.eos_id, attention_dropout=0.0)
    if shape.arch == ARCH_LLAMA:
        from transformers import LlamaConfig
        return LlamaConfig(vocab_size=shape.vocab, hidden_size=shape.d_model, intermediate_size=shape.d_ff,
                           num_hidden_layers=shape.n_layer, num_attention_heads=shape.n_head,
                           num_key_value_heads=shape.n_kv_head, head_dim=shape.head_dim, hidden_act="silu",
                           max_position_embeddings=shape.max_pos, rms_norm_eps=shape.norm_eps, rope_theta=shape.rope_theta,
                           attention_bias=False, mlp_bias=False, tie_word_embeddings=shape.tied_head,
                           bos_token_id=1, eos_token_id=shape.eos_id, attention_dropout=0.0)
    if shape.arch == ARCH_OPT:
        from transformers import OPTConfig
        return OPTConfig(vocab_size=shape.vocab, hidden_size=shape.d_model, num_hidden_layers=shape.n_layer,
                         ffn_dim=shape.d_ff, num_attention_heads=shape.n_head, max_position_embeddings=shape.max_pos,
                         do_layer_norm_before=True, word_embed_proj_dim=shape.d_model, activation_function="relu",
                         dropout=0.0, attention_dropout=0.0, layerdrop=0.0, enable_bias=True,
                         layer_norm_elementwise_affine=True, bos_token_id=shape.eos_id, eos_token_id=shape.eos_id,
                         pad_token_id=1, tie_word_embeddings=True)
    if shape.arch == ARCH_BLOOM:
        from transformers import BloomConfig
        return BloomConfig(vocab_size=shape.vocab, hidden_size=shape.d_model, n_layer=shape.n_layer, n_head=shape.n_head,
                           layer_norm_epsilon=shape.norm_eps, apply_residual_connection_post_layernorm=False,
                           hidden_dropout=0.0, attention_dropout=0.0, bos_token_id=1, eos_token_id=shape.eos_id,
                           pad_token_id=3, tie_word_embeddings=True)
    raise ValueError(shape.arch)


def to_hf_state_dict(shape: DecoderShape, W: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    sd = {}
    if shape.arch == ARCH_FALCON:
        sd["transformer.word_embeddings.weight"] = W["wte"]
        sd["transformer.ln_f.weight"] = W["lnf_g"]
        sd["transformer.ln_f.bias"] = W["lnf_b"]
        sd["lm_head.weight"] = W["lm_head"]
        for l in range(shape.n_layer):
            p, h = f"l{l}.", f"transformer.h.{l}."
            sd[h + "input_layernorm.weight"] = W[p + "ln1_g"]
            sd[h + "input_layernorm.bias"] = W[p + "ln1_b"]
            sd[h + "self_attention.query_key_value.weight"] = W[p + "w_qkv"]
            sd[h + "self_attention.dense.weight"] = W[p + "w_dense"]
            sd[h + "mlp.dense_h_to_4h.weight"] = W[p + "w_up"]
            sd[h + "mlp.dense_4h_to_h.weight"] = W[p + "w_down"]
    elif shape.arch == ARCH_LLAMA:
        # HF:models/llama/modeling_llama.py: separate q/k/v and gate/up projections; the flat layout fuses them by rows
        nq, nk, ff = shape.n_head * shape.head_dim, shape.n_kv_head * shape.head_dim, shape.d_ff
        sd["model.embed_tokens.weight"] = W["wte"]
        sd["model.norm.weight"] = W["lnf_g"]
        sd["lm_head.weight"] = W["lm_head"]
        for l in range(shape.n_layer):
            p, h = f"l{l}.", f"model.layers.{l}."
            sd[h + "input_layernorm.weight"] = W[p + "ln1_g"]
            sd[h + "post_attention_layernorm.weight"] = W[p + "ln2_g"]
            sd[h + "self_attn.q_proj.weight"] = W[p + "w_qkv"][:nq]
            sd[h + "self_attn.k_proj.weight"] = W[p + "w_qkv"][nq:nq + nk]
            sd[h + "self_attn.v_proj.weight"] = W[p + "w_qkv"][nq + nk:]
            if shape.qkv_bias:
                sd[h + "self_attn.q_proj.bias"] = W[p + "b_qkv"][:nq]
                sd[h + "self_attn.k_proj.bias"] = W[p + "b_qkv"][nq:nq + nk]
                sd[h + "self_attn.v_proj.bias"] = W[p + "b_qkv"][nq + nk:]
            sd[h + "self_attn.o_proj.weight"] = W[p + "w_dense"]
            sd[h + "mlp.gate_proj.weight"] = W[p + "w_up"][:ff]
            sd[h + "mlp.up_proj.weight"] = W[p + "w_up"][ff:]
            sd[h + "mlp.down_proj.weight"] = W[p + "w_down"]
    elif shape.arch == ARCH_OPT:
        # HF:models/opt/modeling_opt.py: separate q/k/v projections with biases; fused by rows in the flat layout
        nq = shape.n_head * shape.head_dim
        b = "model.decoder."
        sd[b + "embed_tokens.weight"] = W["wte"]
        sd[b + "embed_positions.weight"] = W["wpe"]
        sd[b + "final_layer_norm.weight"] = W["lnf_g"]
        sd[b + "final_layer_norm.bias"] = W["lnf_b"]
        sd["lm_head.weight"] = W["lm_head"]
        for l in range(shape.n_layer):
            p, h = f"l{l}.", f"{b}layers.{l}."
            sd[h + "self_attn_layer_norm.weight"] = W[p + "ln1_g"]
            sd[h + "self_attn_layer_norm.bias"] = W[p + "ln1_b"]
            sd[h + "final_layer_norm.weight"] = W[p + "ln2_g"]
            sd[h + "final_layer_norm.bias"] = W[p + "ln2_b"]
            for i, n in enumerate(("q_proj", "k_proj", "v_proj")):
                sd[h + f"self_attn.{n}.weight"] = W[p + "w_qkv"][i * nq:(i + 1) * nq]
                sd[h + f"self_attn.{n}.bias"] = W[p + "b_qkv"][i * nq:(i + 1) * nq]
            sd[h + "self_attn.out_proj.weight"] = W[p + "w_dense"]
            sd[h + "self_attn.out_proj.bias"] = W[p + "b_dense"]
            sd[h + "fc1.weight"] = W[p + "w_up"]
            sd[h + "fc1.bias"] = W[p + "b_up"]
            sd[h + "fc2.weight"] = W[p + "w_down"]
            sd[h + "fc2.bias"] = W[p + "b_down"]
    elif shape.arch == ARCH_BLOOM:
        # HF:models/bloom/modeling_bloom.py: fused query_key_value with per-head [q, k, v] rows
        sd["transformer.word_embeddings.weight"] = W["wte"]
        sd["transformer.word_embeddings_layernorm.weight"] = W["emb_ln_g"]
        sd["transformer.word_embeddings_layernorm.bias"] = W["emb_ln_b"]
        sd["transformer.ln_f.weight"] = W["lnf_g"]
        sd["transformer.ln_f.bias"] = W["lnf_b"]
        sd["lm_head.weight"] = W["lm_head"]
        for l in range(shape.n_layer):
            p, h = f"l{l}.", f"transformer.h.{l}."
            sd[h + "input_layernorm.weight"] = W[p + "ln1_g"]
            sd[h + "input_layernorm.bias"] = W[p + "ln1_b"]
            sd[h + "post_attention_layernorm.weight"] = W[p + "ln2_g"]
            sd[h + "post_attention_layernorm.bias"] = W[p + "ln2_b"]
            sd[h + "self_attention.query_key_value.weight"] = W[p + "w_qkv"]
            sd[h + "self_attention.query_key_value.bias"] = W[p + "b_qkv"]
            sd[h + "self_attention.dense.weight"] = W[p + "w_dense"]
            sd[h + "self_attention.dense.bias"] = W[p + "b_dense"]
            sd[h + "mlp.dense_h_to_4h.weight"] = W[p + "w_up"]
            sd[h + "mlp.dense_h_to_4h.bias"] = W[p + "b_up"]
            sd[h + "mlp.dense_4h_to_h.weight"] = W[p + "w_down"]
            sd[h + "mlp.dense_4h_to_h.bias"] = W[p + "b_down"]
    else:
        sd["gpt_neox.embed_in.weight"] = W["wte"]
        sd["gpt_neox.final_layer_norm.weight"] = W["lnf_g"]
        sd["gpt_neox.final_layer_norm.bias"] = W["lnf_b"]
        sd["embed_out.weight"] = W["lm_head"]
        for l in range(shape.n_layer):
            p, h = f"l{l}.", f"gpt_neox.layers.{l}."
            sd[h + "input_layernorm.weight"] = W[p + "ln1_g"]
            sd[h + "input_layernorm.bias"] = W[p + "ln1_b"]
            sd[h + "post_attention_layernorm.weight"] = W[p + "ln2_g"]
            sd[h + "post_attention_layernorm.bias"] = W[p + "ln2_b"]
            sd[h + "attention.query_key_value.weight"] = W[p + "w_qkv"]
            sd[h + "attention.query_key_value.bias"] = W[p + "b_qkv"]
            sd[h + "attention.dense.weight"] = W[p + "w_dense"]
            sd[h + "attention.dense.bias"] = W[p + "b_dense"]
            sd[h + "mlp.dense_h_to_4h.weight"] = W[p + "w_up"]
            sd[h + "mlp.dense_h_to_4h.bias"] = W[p + "b_up"]
            sd[h + "mlp.dense_4h_to_h.weight"] = W[p + "w_down"]
            sd[h + "mlp.dense_4h_to_h.bias"] = W[p + "b_down"]
    return sd


def build_hf_model(shape: DecoderShape, W: Dict[str, torch.Tensor]):
    """Instantiate the HF model on the meta device and bind ``W`` (fp32, CPU) to it; avoids the
    minutes-long HF random init for 7B shapes."""
    import torch.nn as nn
    from transformers import (BloomForCausalLM, FalconForCausalLM, GPTNeoXForCausalLM, LlamaForCausalLM,
                              OPTForCausalLM, Qwen2ForCausalLM)
    cfg = hf_config(shape)
    cls = {ARCH_FALCON: FalconForCausalLM, ARCH_GPT_NEOX: GPTNeoXForCausalLM, ARCH_LLAMA: LlamaForCausalLM,
           ARCH_OPT: OPTForCausalLM, ARCH_BLOOM: BloomForCausalLM}[shape.arch]
    if shape.arch == ARCH_LLAMA and shape.qkv_bias:
        cls = Qwen2ForCausalLM
    with torch.device("meta"):
        model = cls(cfg)
    sd = {k: v.to(torch.float32).contiguous() for k, v in to_hf_state_dict(shape, W).items()}
    missing, unexpected = model.load_state_dict(sd, strict=False, assign=True)
    # rotary inv_freq buffers are non-persistent and were created on meta: rebuild them
    for mod in model.modules():
        if hasattr(mod, "inv_freq") and getattr(mod.inv_freq, "is_meta", False):
            rope_init = getattr(mod, "rope_init_fn", None)
            if rope_init is None:
                rope_init = type(mod).compute_default_rope_parameters
            inv, _ = rope_init(mod.config, torch.device("cpu"))
            mod.register_buffer("inv_freq", inv, persistent=False)
            if hasattr(mod, "original_inv_freq"):
                mod.register_buffer("original_inv_freq", inv.clone(), persistent=False)
    bad = [m for m in missing if "inv_freq" not in m]
    assert not bad, f"missing keys {bad}"
    assert not unexpected, f"unexpected keys {unexpected}"
    still_meta = [n for n, p in list(model.named_parameters()) + list(model.named_buffers()) if p.is_meta]
    assert not still_meta, f"meta tensors left: {still_meta[:4]}"
    model.eval()
    return model


# ---- encoder-decoder (T5 family) ------------------------------------------------------------------
def t5_config(shape):
    """HF T5Config equivalent to an ``EncDecShape`` (llm_interpretation_replication_b200.t5)."""
    from transformers import T5Config
    return T5Config(vocab_size=shape.vocab, d_model=shape.d_model, d_kv=shape.head_dim, d_ff=shape.d_ff,
                    num_layers=shape.n_enc_layer, num_decoder_layers=shape.n_dec_layer, num_heads=shape.n_head,
                    relative_attention_num_buckets=shape.rel_buckets, relative_attention_max_distance=shape.rel_max_distance,
                    dropout_rate=0.0, layer_norm_epsilon=shape.norm_eps,
                    feed_forward_proj="gated-gelu" if shape.gated else "relu", is_encoder_decoder=True, use_cache=True,
                    pad_token_id=shape.pad_id, eos_token_id=shape.eos_id, decoder_start_token_id=shape.decoder_start_id,
                    tie_word_embeddings=shape.tied_head)


def t5_state_dict(shape, W: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Flat T5 layout -> HF names (HF:models/t5/modeling_t5.py: separate q/k/v/o per attention, wi or wi_0/wi_1 + wo, the
    relative bias on block 0 of each stack)."""
    inner, ff = shape.inner, shape.d_ff
    sd = {"shared.weight": W["wte"], "encoder.embed_tokens.weight": W["wte"], "decoder.embed_tokens.weight": W["wte"],
          "encoder.final_layer_norm.weight": W["enc_lnf_g"], "decoder.final_layer_norm.weight": W["dec_lnf_g"],
          "lm_head.weight": W["lm_head"],
          "encoder.block.0.layer.0.SelfAttention.relative_attention_bias.weight": W["enc_rel_bias"],
          "decoder.block.0.layer.0.SelfAttention.relative_attention_bias.weight": W["dec_rel_bias"]}

    def ffn(h, p):
        if shape.gated:
            sd[h + "DenseReluDense.wi_0.weight"] = W[p + "w_up"][:ff]
            sd[h + "DenseReluDense.wi_1.weight"] = W[p + "w_up"][ff:]
        else:
            sd[h + "DenseReluDense.wi.weight"] = W[p + "w_up"]
        sd[h + "DenseReluDense.wo.weight"] = W[p + "w_down"]

    for l in range(shape.n_enc_layer):
        p, h = f"e{l}.", f"encoder.block.{l}.layer."
        sd[h + "0.layer_norm.weight"] = W[p + "ln1_g"]
        for i, n in enumerate("qkv"):
            sd[h + f"0.SelfAttention.{n}.weight"] = W[p + "w_qkv"][i * inner:(i + 1) * inner]
        sd[h + "0.SelfAttention.o.weight"] = W[p + "w_o"]
        sd[h + "1.layer_norm.weight"] = W[p + "ln2_g"]
        ffn(h + "1.", p)
    for l in range(shape.n_dec_layer):
        p, h = f"d{l}.", f"decoder.block.{l}.layer."
        sd[h + "0.layer_norm.weight"] = W[p + "ln1_g"]
        for i, n in enumerate("qkv"):
            sd[h + f"0.SelfAttention.{n}.weight"] = W[p + "w_qkv"][i * inner:(i + 1) * inner]
        sd[h + "0.SelfAttention.o.weight"] = W[p + "w_o"]
        sd[h + "1.layer_norm.weight"] = W[p + "ln2_g"]
        sd[h + "1.EncDecAttention.q.weight"] = W[p + "w_cq"]
        sd[h + "1.EncDecAttention.k.weight"] = W[p + "w_ckv"][:inner]
        sd[h + "1.EncDecAttention.v.weight"] = W[p + "w_ckv"][inner:]
        sd[h + "1.EncDecAttention.o.weight"] = W[p + "w_co"]
        sd[h + "2.layer_norm.weight"] = W[p + "ln3_g"]
        ffn(h + "2.", p)
    return sd


def build_hf_t5(shape, W: Dict[str, torch.Tensor]):
    """``T5ForConditionalGeneration`` (what ``AutoModelForSeq2SeqLM.from_pretrained`` returns for the reference's
    encoder-decoder models, compare_instruct_models.py:470-476) holding ``W`` (fp32, CPU)."""
    from transformers import T5ForConditionalGeneration
    with torch.device("meta"):
        model = T5ForConditionalGeneration(t5_config(shape))
    sd = {k: v.to(torch.float32).contiguous() for k, v in t5_state_dict(shape, W).items()}
    missing, unexpected = model.load_state_dict(sd, strict=False, assign=True)
    assert not missing, f"missing keys {missing}"
    assert not unexpected, f"unexpected keys {unexpected}"
    still_meta = [n for n, p in list(model.named_parameters()) + list(model.named_buffers()) if p.is_meta]
    assert not still_meta, f"meta tensors left: {still_meta[:4]}"
    model.eval()
    return model
