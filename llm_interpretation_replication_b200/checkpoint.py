"""Local Hugging Face checkpoints and tokenizers -> the scorer.

Replaces ``AutoTokenizer.from_pretrained`` / ``AutoModelForCausalLM.from_pretrained`` of the reference
(``analysis/run_base_vs_instruct_100q.py:416-450``, ``compare_instruct_models.py:178-226``) for checkpoints that are
already on disk (there is no hub access here): a directory with ``config.json`` plus ``*.safetensors`` (single file or
sharded with ``model.safetensors.index.json``) or ``pytorch_model*.bin``.

* ``shape_from_hf_config``  (also the Llama family, OPT, BLOOM, Qwen2, and -- key names from their published custom code, which
  this image cannot instantiate -- the first-generation Qwen and Baichuan2-7B)  Falcon (``model_type`` "falcon" or the original "RefinedWebModel" key names; multi-query,
  parallel attention, rotary: HF:models/falcon/configuration_falcon.py:53-76) and GPT-NeoX (StableLM-alpha, RedPajama,
  Pythia/Dolly, h2oGPT: HF:models/gpt_neox/configuration_gpt_neox.py) configs to a ``DecoderShape``; anything else
  raises ``KeyError``, which the entry points turn into the reference's NaN rows for a model that fails to load.
* ``flat_from_hf``          inverse of ``weights.to_hf_state_dict``.  The fused QKV matrices keep HF's own row order
  (Falcon ``[q heads..., k, v]`` HF falcon:285-288; NeoX per-head ``[q, k, v]`` HF gpt_neox:212-215): the kernels
  index them through ``qkv_layout``, nothing is permuted at load time.
* ``t5_shape_from_hf_config`` / ``t5_flat_from_hf``  the same for the encoder-decoder checkpoints of the reference's lists
  (T5 v1.1, Flan-T5, T0, Tk-Instruct) -> ``t5.EncDecShape`` and the layout of ``t5.slot_names``.
* ``load_checkpoint``       streams tensor by tensor to the device in the compute dtype (a 7B model never holds a
  second host copy) and returns ``(shape, weights)`` ready for ``Scorer``.
* ``load_tokenizer``        ``AutoTokenizer.from_pretrained(dir, local_files_only=True)`` passed through unchanged, so
  the F4 quirk (`` Yes`` looked up through ``tokenizer(...).input_ids[0]`` with or without special tokens) behaves
  exactly as in the reference; ``pad_token = eos_token`` when the tokenizer has none (reference :420-421).
"""
from __future__ import annotations

import glob
import json
import os
import re
from typing import Dict, Iterator, Optional, Tuple

import torch

from .shapes import (ACT_GELU_TANH, ACT_RELU, ARCH_BLOOM, ARCH_FALCON, ARCH_GPT_NEOX, ARCH_LLAMA, ARCH_OPT, MLP_SWIGLU, NORM_RMS,
                     POS_ALIBI, POS_LEARNED, QKV_FALCON_MQA, QKV_NEOX_INTERLEAVED, DecoderShape)
from . import weights as _weights


def _rope(cfg: dict, theta_keys, pct_default):
    """(rope_theta, partial rotary factor) from either generation of config keys: the hub checkpoints of the sweep carry
    ``rotary_emb_base`` / ``rotary_pct`` (GPT-NeoX) or ``rope_theta`` (Falcon); transformers >= 5 re-serialises them as
    ``rope_parameters = {rope_theta, partial_rotary_factor, rope_type}``."""
    rp = cfg.get("rope_parameters") or {}
    if rp.get("rope_type", "default") not in ("default", None):
        raise KeyError(f"rope_type {rp.get('rope_type')!r} is not supported")
    if cfg.get("rope_scaling") and (cfg["rope_scaling"].get("rope_type") or cfg["rope_scaling"].get("type")) not in (None, "default"):
        raise KeyError("rope_scaling is not supported")
    theta = rp.get("rope_theta")
    for k in theta_keys:
        if theta is None:
            theta = cfg.get(k)
    pct = rp.get("partial_rotary_factor", cfg.get("rotary_pct", cfg.get("partial_rotary_factor", pct_default)))
    return float(10000.0 if theta is None else theta), float(pct)


def shape_from_hf_config(cfg: dict, name: str = "checkpoint") -> DecoderShape:
    mt = cfg.get("model_type", "")
    if mt in ("falcon", "RefinedWebModel", "RefinedWeb"):
        n_layer = cfg.get("num_hidden_layers", cfg.get("n_layer"))
        n_head = cfg.get("num_attention_heads", cfg.get("n_head"))
        d = cfg["hidden_size"]
        if cfg.get("new_decoder_architecture", False) or mt == "RefinedWeb":
            raise KeyError(f"{name}: Falcon new_decoder_architecture (40B/180B layout) is not supported")
        if not cfg.get("multi_query", True) or not cfg.get("parallel_attn", True):
            raise KeyError(f"{name}: only multi-query, parallel-attention Falcon is supported")
        if cfg.get("alibi", False) or cfg.get("bias", False):
            raise KeyError(f"{name}: Falcon with alibi or biases is not supported")
        hd = d // n_head
        theta, _ = _rope(cfg, ("rope_theta",), 1.0)
        return DecoderShape(name, ARCH_FALCON, int(n_layer), int(d), int(n_head), 1, hd, int(cfg.get("ffn_hidden_size") or 4 * d),
                            int(cfg["vocab_size"]), hd, rope_theta=theta,
                            norm_eps=float(cfg.get("layer_norm_epsilon", 1e-5)), parallel_residual=True, n_ln_per_layer=1,
                            has_bias=False, tied_head=True, qkv_layout=QKV_FALCON_MQA,
                            eos_id=int(cfg.get("eos_token_id", 11)), max_pos=int(cfg.get("max_position_embeddings", 2048)))
    if mt == "gpt_neox":
        if cfg.get("hidden_act", "gelu") != "gelu":
            raise KeyError(f"{name}: GPT-NeoX hidden_act {cfg.get('hidden_act')!r} is not supported (erf GELU only)")
        if not cfg.get("attention_bias", True):
            raise KeyError(f"{name}: GPT-NeoX without attention biases is not supported")
        d, n_head = int(cfg["hidden_size"]), int(cfg["num_attention_heads"])
        hd = d // n_head
        theta, pct = _rope(cfg, ("rotary_emb_base", "rope_theta"), 0.25)
        rd = int(hd * pct)
        eos = cfg.get("eos_token_id", 0)
        eos = eos[0] if isinstance(eos, (list, tuple)) else eos
        return DecoderShape(name, ARCH_GPT_NEOX, int(cfg["num_hidden_layers"]), d, n_head, n_head, hd,
                            int(cfg["intermediate_size"]), int(cfg["vocab_size"]), rd,
                            rope_theta=theta,
                            norm_eps=float(cfg.get("layer_norm_eps", 1e-5)),
                            parallel_residual=bool(cfg.get("use_parallel_residual", True)), n_ln_per_layer=2, has_bias=True,
                            tied_head=bool(cfg.get("tie_word_embeddings", False)), qkv_layout=QKV_NEOX_INTERLEAVED,
                            eos_id=int(eos), max_pos=int(cfg.get("max_position_embeddings", 2048)))
    if mt in ("llama", "mistral"):
        # HF:models/llama/configuration_llama.py, models/mistral/configuration_mistral.py (Mistral-7B-Instruct-v0.3 of the
        # instruct sweep, compare_instruct_models.py:157).  A sliding window shorter than the prompts would change the
        # attention mask and is refused; v0.3 has none.
        if cfg.get("hidden_act", "silu") != "silu":
            raise KeyError(f"{name}: hidden_act {cfg.get('hidden_act')!r} is not supported (SiLU-gated MLP only)")
        if cfg.get("attention_bias", False) or cfg.get("mlp_bias", False):
            raise KeyError(f"{name}: Llama-family checkpoints with projection biases are not supported")
        max_pos = int(cfg.get("max_position_embeddings", 4096))
        sw = cfg.get("sliding_window")
        if sw is not None and int(sw) < 4096:
            raise KeyError(f"{name}: sliding_window {sw} is not supported")
        d, n_head = int(cfg["hidden_size"]), int(cfg["num_attention_heads"])
        hd = int(cfg.get("head_dim") or d // n_head)
        theta, pct = _rope(cfg, ("rope_theta",), 1.0)
        if pct != 1.0:
            raise KeyError(f"{name}: partial rotary in a Llama-family model is not supported")
        eos = cfg.get("eos_token_id", 2)
        eos = eos[0] if isinstance(eos, (list, tuple)) else eos
        return DecoderShape(name, ARCH_LLAMA, int(cfg["num_hidden_layers"]), d, n_head,
                            int(cfg.get("num_key_value_heads") or n_head), hd, int(cfg["intermediate_size"]),
                            int(cfg["vocab_size"]), hd, rope_theta=theta, norm_eps=float(cfg.get("rms_norm_eps", 1e-6)),
                            parallel_residual=False, n_ln_per_layer=2, has_bias=False,
                            tied_head=bool(cfg.get("tie_word_embeddings", False)), qkv_layout=QKV_FALCON_MQA, eos_id=int(eos),
                            max_pos=min(max_pos, 32768), norm_type=NORM_RMS, mlp_type=MLP_SWIGLU)
    if mt == "qwen2":
        # HF:models/qwen2/configuration_qwen2.py: the Llama block with q/k/v biases (the in-library successor of the reference's Qwen-1)
        if cfg.get("hidden_act", "silu") != "silu":
            raise KeyError(f"{name}: hidden_act {cfg.get('hidden_act')!r} is not supported (SiLU-gated MLP only)")
        if cfg.get("use_sliding_window", False):
            raise KeyError(f"{name}: Qwen2 with a sliding window is not supported")
        d, n_head = int(cfg["hidden_size"]), int(cfg["num_attention_heads"])
        theta, pct = _rope(cfg, ("rope_theta",), 1.0)
        eos = cfg.get("eos_token_id", 151643)
        eos = eos[0] if isinstance(eos, (list, tuple)) else eos
        return DecoderShape(name, ARCH_LLAMA, int(cfg["num_hidden_layers"]), d, n_head, int(cfg.get("num_key_value_heads") or n_head),
                            int(cfg.get("head_dim") or d // n_head), int(cfg["intermediate_size"]), int(cfg["vocab_size"]),
                            int(cfg.get("head_dim") or d // n_head), rope_theta=theta, norm_eps=float(cfg.get("rms_norm_eps", 1e-6)),
                            parallel_residual=False, n_ln_per_layer=2, has_bias=False, tied_head=bool(cfg.get("tie_word_embeddings", False)),
                            qkv_layout=QKV_FALCON_MQA, eos_id=int(eos), max_pos=min(int(cfg.get("max_position_embeddings", 32768)), 32768),
                            norm_type=NORM_RMS, mlp_type=MLP_SWIGLU, qkv_bias=True)
    if mt == "qwen":
        # Qwen/Qwen-7B(-Chat) (compare_instruct_models.py:159): custom code on the hub (QWenLMHeadModel), so the key names come from
        # its published config.json / modeling_qwen.py, not from a class this image can instantiate -- parity of the MATH is
        # pinned through Qwen2 (same block), the key mapping is unpinned.  intermediate_size counts w1 and w2 together.
        if not cfg.get("no_bias", True):
            raise KeyError(f"{name}: Qwen with biases beyond c_attn is not supported")
        if float(cfg.get("rotary_pct", 1.0)) != 1.0:
            raise KeyError(f"{name}: partial rotary in Qwen is not supported")
        d, n_head = int(cfg["hidden_size"]), int(cfg["num_attention_heads"])
        hd = int(cfg.get("kv_channels") or d // n_head)
        return DecoderShape(name, ARCH_LLAMA, int(cfg["num_hidden_layers"]), d, n_head, n_head, hd, int(cfg["intermediate_size"]) // 2,
                            int(cfg["vocab_size"]), hd, rope_theta=float(cfg.get("rotary_emb_base", 10000.0)),
                            norm_eps=float(cfg.get("layer_norm_epsilon", 1e-6)), parallel_residual=False, n_ln_per_layer=2, has_bias=False,
                            tied_head=bool(cfg.get("tie_word_embeddings", False)), qkv_layout=QKV_FALCON_MQA,
                            eos_id=int(cfg.get("eos_token_id", 151643)), max_pos=min(int(cfg.get("seq_length", 8192)), 32768),
                            norm_type=NORM_RMS, mlp_type=MLP_SWIGLU, qkv_bias=True)
    if mt == "baichuan":
        # baichuan-inc/Baichuan2-7B-Chat (compare_instruct_models.py:147): custom code on the hub (BaichuanForCausalLM); Llama block
        # with q|k|v packed in W_pack.  The 13B models use ALiBi instead of rotary positions and are refused.
        if int(cfg["num_hidden_layers"]) >= 40 or int(cfg["hidden_size"]) > 4096:
            raise KeyError(f"{name}: Baichuan 13B (ALiBi) is not supported")
        d, n_head = int(cfg["hidden_size"]), int(cfg["num_attention_heads"])
        return DecoderShape(name, ARCH_LLAMA, int(cfg["num_hidden_layers"]), d, n_head, n_head, d // n_head, int(cfg["intermediate_size"]),
                            int(cfg["vocab_size"]), d // n_head, rope_theta=10000.0, norm_eps=float(cfg.get("rms_norm_eps", 1e-6)),
                            parallel_residual=False, n_ln_per_layer=2, has_bias=False, tied_head=False, qkv_layout=QKV_FALCON_MQA,
                            eos_id=int(cfg.get("eos_token_id", 2)), max_pos=int(cfg.get("model_max_length", cfg.get("max_position_embeddings", 4096))),
                            norm_type=NORM_RMS, mlp_type=MLP_SWIGLU)
    if mt == "opt":
        # HF:models/opt/configuration_opt.py (facebook/opt-iml-1.3b of the instruct sweep, compare_instruct_models.py:151)
        d = int(cfg["hidden_size"])
        if int(cfg.get("word_embed_proj_dim", d)) != d or not cfg.get("do_layer_norm_before", True):
            raise KeyError(f"{name}: OPT with projected embeddings or post-LayerNorm (opt-350m) is not supported")
        if cfg.get("activation_function", "relu") != "relu" or not cfg.get("enable_bias", True):
            raise KeyError(f"{name}: OPT variant (activation / no biases) is not supported")
        n_head = int(cfg["num_attention_heads"])
        return DecoderShape(name, ARCH_OPT, int(cfg["num_hidden_layers"]), d, n_head, n_head, d // n_head, int(cfg["ffn_dim"]),
                            int(cfg["vocab_size"]), 0, norm_eps=1e-5, parallel_residual=False, n_ln_per_layer=2, has_bias=True,
                            tied_head=bool(cfg.get("tie_word_embeddings", True)), qkv_layout=QKV_FALCON_MQA,
                            eos_id=int(cfg.get("eos_token_id", 2)), max_pos=int(cfg.get("max_position_embeddings", 2048)),
                            pos_type=POS_LEARNED, pos_offset=2, act_type=ACT_RELU)
    if mt == "bloom":
        # HF:models/bloom/configuration_bloom.py (bigscience/bloomz-7b1, compare_instruct_models.py:148)
        if cfg.get("apply_residual_connection_post_layernorm", False):
            raise KeyError(f"{name}: BLOOM with apply_residual_connection_post_layernorm is not supported")
        d = int(cfg.get("hidden_size", cfg.get("n_embed")))
        n_head = int(cfg.get("n_head", cfg.get("num_attention_heads")))
        return DecoderShape(name, ARCH_BLOOM, int(cfg.get("n_layer", cfg.get("num_hidden_layers"))), d, n_head, n_head,
                            d // n_head, 4 * d, int(cfg["vocab_size"]), 0, norm_eps=float(cfg.get("layer_norm_epsilon", 1e-5)),
                            parallel_residual=False, n_ln_per_layer=2, has_bias=True, tied_head=True,
                            qkv_layout=QKV_NEOX_INTERLEAVED, eos_id=int(cfg.get("eos_token_id", 2)),
                            max_pos=int(cfg.get("seq_length", 2048)), pos_type=POS_ALIBI, act_type=ACT_GELU_TANH, emb_norm=True)
    raise KeyError(f"{name}: architecture {mt!r} is not supported by the B200 scorer")


_FALCON_GLOBAL = {"transformer.word_embeddings.weight": "wte", "transformer.ln_f.weight": "lnf_g",
                  "transformer.ln_f.bias": "lnf_b", "lm_head.weight": "lm_head"}
_FALCON_LAYER = {"input_layernorm.weight": "ln1_g", "input_layernorm.bias": "ln1_b",
                 "self_attention.query_key_value.weight": "w_qkv", "self_attention.dense.weight": "w_dense",
                 "mlp.dense_h_to_4h.weight": "w_up", "mlp.dense_4h_to_h.weight": "w_down"}
_NEOX_GLOBAL = {"gpt_neox.embed_in.weight": "wte", "gpt_neox.final_layer_norm.weight": "lnf_g",
                "gpt_neox.final_layer_norm.bias": "lnf_b", "embed_out.weight": "lm_head"}
_NEOX_LAYER = {"input_layernorm.weight": "ln1_g", "input_layernorm.bias": "ln1_b",
               "post_attention_layernorm.weight": "ln2_g", "post_attention_layernorm.bias": "ln2_b",
               "attention.query_key_value.weight": "w_qkv", "attention.query_key_value.bias": "b_qkv",
               "attention.dense.weight": "w_dense", "attention.dense.bias": "b_dense",
               "mlp.dense_h_to_4h.weight": "w_up", "mlp.dense_h_to_4h.bias": "b_up",
               "mlp.dense_4h_to_h.weight": "w_down", "mlp.dense_4h_to_h.bias": "b_down"}
_LLAMA_GLOBAL = {"model.embed_tokens.weight": "wte", "model.norm.weight": "lnf_g", "lm_head.weight": "lm_head",
                 "transformer.wte.weight": "wte", "transformer.ln_f.weight": "lnf_g"}            # (Qwen-1 names)
# q/k/v and gate/up arrive as separate tensors and are fused by rows (":0" = first block of the fused matrix, ...)
_LLAMA_LAYER = {"input_layernorm.weight": "ln1_g", "post_attention_layernorm.weight": "ln2_g",
                "self_attn.q_proj.weight": "w_qkv:0", "self_attn.k_proj.weight": "w_qkv:1", "self_attn.v_proj.weight": "w_qkv:2",
                "self_attn.o_proj.weight": "w_dense", "mlp.gate_proj.weight": "w_up:0", "mlp.up_proj.weight": "w_up:1",
                "mlp.down_proj.weight": "w_down",
                # Qwen2: biases on q, k, v
                "self_attn.q_proj.bias": "b_qkv:0", "self_attn.k_proj.bias": "b_qkv:1", "self_attn.v_proj.bias": "b_qkv:2",
                # Baichuan2: q|k|v packed by rows (modeling_baichuan.py: W_pack(x).unflatten(-1, (3, hidden)))
                "self_attn.W_pack.weight": "w_qkv",
                # Qwen-1 (modeling_qwen.py): c_attn = q|k|v with bias; MLP = c_proj(w1(x) * silu(w2(x))): w2 is the gate
                "ln_1.weight": "ln1_g", "ln_2.weight": "ln2_g", "attn.c_attn.weight": "w_qkv", "attn.c_attn.bias": "b_qkv",
                "attn.c_proj.weight": "w_dense", "mlp.w2.weight": "w_up:0", "mlp.w1.weight": "w_up:1", "mlp.c_proj.weight": "w_down"}
_OPT_GLOBAL = {"model.decoder.embed_tokens.weight": "wte", "model.decoder.embed_positions.weight": "wpe",
               "model.decoder.final_layer_norm.weight": "lnf_g", "model.decoder.final_layer_norm.bias": "lnf_b",
               "lm_head.weight": "lm_head"}
_OPT_LAYER = {"self_attn_layer_norm.weight": "ln1_g", "self_attn_layer_norm.bias": "ln1_b", "final_layer_norm.weight": "ln2_g",
              "final_layer_norm.bias": "ln2_b", "self_attn.q_proj.weight": "w_qkv:0", "self_attn.k_proj.weight": "w_qkv:1",
              "self_attn.v_proj.weight": "w_qkv:2", "self_attn.q_proj.bias": "b_qkv:0", "self_attn.k_proj.bias": "b_qkv:1",
              "self_attn.v_proj.bias": "b_qkv:2", "self_attn.out_proj.weight": "w_dense", "self_attn.out_proj.bias": "b_dense",
              "fc1.weight": "w_up", "fc1.bias": "b_up", "fc2.weight": "w_down", "fc2.bias": "b_down"}
_BLOOM_GLOBAL = {"transformer.word_embeddings.weight": "wte", "transformer.word_embeddings_layernorm.weight": "emb_ln_g",
                 "transformer.word_embeddings_layernorm.bias": "emb_ln_b", "transformer.ln_f.weight": "lnf_g",
                 "transformer.ln_f.bias": "lnf_b", "lm_head.weight": "lm_head"}
_BLOOM_LAYER = {"input_layernorm.weight": "ln1_g", "input_layernorm.bias": "ln1_b", "post_attention_layernorm.weight": "ln2_g",
                "post_attention_layernorm.bias": "ln2_b", "self_attention.query_key_value.weight": "w_qkv",
                "self_attention.query_key_value.bias": "b_qkv", "self_attention.dense.weight": "w_dense",
                "self_attention.dense.bias": "b_dense", "mlp.dense_h_to_4h.weight": "w_up", "mlp.dense_h_to_4h.bias": "b_up",
                "mlp.dense_4h_to_h.weight": "w_down", "mlp.dense_4h_to_h.bias": "b_down"}
_LAYER_RE = re.compile(r"^(?:transformer\.h|gpt_neox\.layers|model\.layers|model\.decoder\.layers)\.(\d+)\.(.+)$")


def flat_name(shape: DecoderShape, hf_name: str) -> Optional[str]:
    """Name of an HF state-dict entry in the flat layout, or None for entries the scorer does not use (rotary
    ``inv_freq`` buffers, causal-mask buffers)."""
    glob_map, layer_map = {ARCH_FALCON: (_FALCON_GLOBAL, _FALCON_LAYER), ARCH_GPT_NEOX: (_NEOX_GLOBAL, _NEOX_LAYER),
                           ARCH_LLAMA: (_LLAMA_GLOBAL, _LLAMA_LAYER), ARCH_OPT: (_OPT_GLOBAL, _OPT_LAYER),
                           ARCH_BLOOM: (_BLOOM_GLOBAL, _BLOOM_LAYER)}[shape.arch]
    if hf_name in glob_map:
        return glob_map[hf_name]
    m = _LAYER_RE.match(hf_name)
    if m and m.group(2) in layer_map:
        return f"l{int(m.group(1))}.{layer_map[m.group(2)]}"
    return None


def expected_names(shape: DecoderShape):
    names = [n for n in _weights.slot_names(shape) if n is not None and not n.startswith("rope_") and n != "alibi_slopes"]
    out = []
    for n in names:
        leaf = n.split(".")[-1]
        if leaf in ("ln2_g", "ln2_b") and shape.n_ln_per_layer != 2:
            continue
        if leaf.startswith("b_") and not shape.has_bias and not (leaf == "b_qkv" and shape.qkv_bias):
            continue
        if leaf in ("lnf_b", "ln1_b", "ln2_b") and shape.norm_type == NORM_RMS:
            continue
        if (leaf == "wpe" and shape.pos_type != POS_LEARNED) or (leaf in ("emb_ln_g", "emb_ln_b") and not shape.emb_norm):
            continue
        out.append(n)
    return out


def _check_shape(shape: DecoderShape, name: str, t: torch.Tensor):
    d, ff, V = shape.d_model, shape.d_ff, shape.vocab
    leaf = name.split(".")[-1]
    nq, nk = shape.n_head * shape.head_dim, shape.n_kv_head * shape.head_dim
    part = {"w_qkv:0": (nq, d), "w_qkv:1": (nk, d), "w_qkv:2": (nk, d), "w_up:0": (ff, d), "w_up:1": (ff, d),
            "b_qkv:0": (nq,), "b_qkv:1": (nk,), "b_qkv:2": (nk,)}
    if leaf in part:
        if tuple(t.shape) != part[leaf]:
            raise ValueError(f"checkpoint tensor {name} has shape {tuple(t.shape)}, the config implies {part[leaf]}")
        return
    want = {"wte": (V, d), "wpe": (shape.max_pos + shape.pos_offset, d), "emb_ln_g": (d,), "emb_ln_b": (d,), "lm_head": (V, d), "lnf_g": (d,), "lnf_b": (d,), "ln1_g": (d,), "ln1_b": (d,), "ln2_g": (d,),
            "ln2_b": (d,), "w_qkv": (shape.qkv_dim, d), "b_qkv": (shape.qkv_dim,), "w_dense": (d, shape.n_head * shape.head_dim),
            "b_dense": (d,), "w_up": (shape.up_dim, d), "b_up": (ff,), "w_down": (d, ff), "b_down": (d,)}[leaf]
    if tuple(t.shape) != want:
        raise ValueError(f"checkpoint tensor {name} has shape {tuple(t.shape)}, the config implies {want}")


def flat_from_hf(shape: DecoderShape, tensors, device="cpu", weight_dtype=torch.float32, norm_head: bool = False) -> Dict[str, torch.Tensor]:
    """``tensors``: a state dict or an iterator of (hf_name, tensor).  Matrices are converted to ``weight_dtype``,
    vectors to fp32, each moved to ``device`` as it arrives.  ``norm_head``: the rows of lm_head are L2-normalised once
    (Baichuan2's NormHead does so on every inference call, modeling_baichuan.py NormHead.forward)."""
    items = tensors.items() if hasattr(tensors, "items") else tensors
    W: Dict[str, torch.Tensor] = {}
    for hf_name, t in items:
        n = flat_name(shape, hf_name)
        if n is None:
            continue
        _check_shape(shape, n, t)
        W[n] = t.to(device=device, dtype=weight_dtype if _weights.is_matrix(n.split(":")[0]) else torch.float32).contiguous()
        base = n.split(":")[0]
        if base != n:                                  # a block of a fused matrix: fuse once every block is here
            blocks = [f"{base}:{i}" for i in range(3 if base.endswith("_qkv") else 2)]
            if all(b in W for b in blocks):
                W[base] = torch.cat([W.pop(b) for b in blocks], dim=0).contiguous()
    if "lm_head" not in W:
        if not shape.tied_head:
            raise ValueError("checkpoint has no output embedding and the config does not tie it to the input embedding")
        W["lm_head"] = W["wte"]
    if norm_head:
        W["lm_head"] = torch.nn.functional.normalize(W["lm_head"].float(), dim=-1).to(W["lm_head"].dtype).contiguous()
    if shape.pos_type == POS_ALIBI:                    # not stored in checkpoints: HF recomputes the slopes from n_head
        W["alibi_slopes"] = _weights.alibi_slopes(shape.n_head).to(device)
    missing = [n for n in expected_names(shape) if n not in W]
    if missing:
        raise ValueError(f"checkpoint is missing {len(missing)} tensors, e.g. {missing[:4]}")
    return W


# ---- encoder-decoder (T5 family) ------------------------------------------------------------------------------------
def t5_shape_from_hf_config(cfg: dict, name: str = "checkpoint"):
    """HF:models/t5/configuration_t5.py -> ``t5.EncDecShape`` (google/t5-v1_1-base, google/flan-t5-base, bigscience/T0_3B,
    allenai/tk-instruct-3b-def: compare_base_vs_instruct.py:137-138, compare_instruct_models.py:147,150)."""
    from .t5 import EncDecShape
    ffp = cfg.get("feed_forward_proj", "relu")
    if ffp not in ("relu", "gated-gelu"):
        raise KeyError(f"{name}: T5 feed_forward_proj {ffp!r} is not supported (relu, gated-gelu)")
    if int(cfg.get("d_kv", 64)) != 64:
        raise KeyError(f"{name}: T5 d_kv {cfg.get('d_kv')} is not supported (64)")
    # Original T5 ties the head to the embedding and scales the decoder output by d_model^-0.5; v1.1 does neither.  Hub
    # configs say so through `tie_word_embeddings`; configs re-saved by transformers >= 5 through `scale_decoder_outputs`
    # (HF:models/t5/configuration_t5.py:77-83).  Whether a separate lm_head exists is read off the checkpoint itself.
    scaled = bool(cfg["scale_decoder_outputs"]) if "scale_decoder_outputs" in cfg else cfg.get("tie_word_embeddings", True) is not False
    n_enc = int(cfg["num_layers"])
    n_dec = cfg.get("num_decoder_layers")
    eos = cfg.get("eos_token_id", 1)
    eos = eos[0] if isinstance(eos, (list, tuple)) else eos
    return EncDecShape(name, n_enc, int(n_dec if n_dec is not None else n_enc), int(cfg["d_model"]), int(cfg["num_heads"]),
                       int(cfg["d_ff"]), int(cfg["vocab_size"]), head_dim=64, norm_eps=float(cfg.get("layer_norm_epsilon", 1e-6)),
                       gated=ffp == "gated-gelu", rel_buckets=int(cfg.get("relative_attention_num_buckets", 32)),
                       rel_max_distance=int(cfg.get("relative_attention_max_distance", 128)),
                       tied_head=scaled, eos_id=int(eos), pad_id=int(cfg.get("pad_token_id", 0)),
                       decoder_start_id=int(cfg.get("decoder_start_token_id", 0)), max_pos=int(cfg.get("n_positions", 512)))


_T5_GLOBAL = {"shared.weight": "wte", "encoder.final_layer_norm.weight": "enc_lnf_g", "decoder.final_layer_norm.weight": "dec_lnf_g",
              "lm_head.weight": "lm_head",
              "encoder.block.0.layer.0.SelfAttention.relative_attention_bias.weight": "enc_rel_bias",
              "decoder.block.0.layer.0.SelfAttention.relative_attention_bias.weight": "dec_rel_bias"}
_T5_ATT = {"q.weight": "w_qkv:0", "k.weight": "w_qkv:1", "v.weight": "w_qkv:2", "o.weight": "w_o"}
_T5_CROSS = {"q.weight": "w_cq", "k.weight": "w_ckv:0", "v.weight": "w_ckv:1", "o.weight": "w_co"}
_T5_FF = {"wi.weight": "w_up", "wi_0.weight": "w_up:0", "wi_1.weight": "w_up:1", "wo.weight": "w_down"}
_T5_RE = re.compile(r"^(encoder|decoder)\.block\.(\d+)\.layer\.(\d)\.(layer_norm\.weight|SelfAttention\.(.+)|EncDecAttention\.(.+)|DenseReluDense\.(.+))$")


def t5_flat_name(hf_name: str) -> Optional[str]:
    if hf_name in _T5_GLOBAL:
        return _T5_GLOBAL[hf_name]
    m = _T5_RE.match(hf_name)
    if not m:
        return None      # encoder/decoder.embed_tokens.weight (copies of shared.weight) and anything else unused
    stack, l, sub = m.group(1), int(m.group(2)), int(m.group(3))
    p = f"{'e' if stack == 'encoder' else 'd'}{l}."
    ff_sub = 1 if stack == "encoder" else 2
    if m.group(4) == "layer_norm.weight":
        return p + ("ln1_g" if sub == 0 else "ln2_g" if (sub == 1) else "ln3_g")
    if m.group(5) is not None and sub == 0:
        return p + _T5_ATT[m.group(5)] if m.group(5) in _T5_ATT else None
    if m.group(6) is not None and stack == "decoder" and sub == 1:
        return p + _T5_CROSS[m.group(6)] if m.group(6) in _T5_CROSS else None
    if m.group(7) is not None and sub == ff_sub:
        return p + _T5_FF[m.group(7)] if m.group(7) in _T5_FF else None
    return None


def t5_flat_from_hf(shape, tensors, device="cpu", weight_dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """HF T5 state dict (or an iterator of (name, tensor)) -> the flat layout of ``t5.slot_names`` (q/k/v, cross k/v and
    wi_0/wi_1 fused by rows as they arrive).  The relative-bias tables stay in checkpoint form ([buckets, heads]);
    ``EncDecScorer`` expands them."""
    from . import t5 as _t5
    items = tensors.items() if hasattr(tensors, "items") else tensors
    d, ff, V, inner = shape.d_model, shape.d_ff, shape.vocab, shape.inner
    want = {"wte": (V, d), "lm_head": (V, d), "enc_lnf_g": (d,), "dec_lnf_g": (d,), "enc_rel_bias": (shape.rel_buckets, shape.n_head),
            "dec_rel_bias": (shape.rel_buckets, shape.n_head), "ln1_g": (d,), "ln2_g": (d,), "ln3_g": (d,), "w_o": (d, inner),
            "w_co": (d, inner), "w_cq": (inner, d), "w_down": (d, ff), "w_up": (ff, d)}
    W: Dict[str, torch.Tensor] = {}
    for hf_name, t in items:
        n = t5_flat_name(hf_name)
        if n is None:
            continue
        leaf = n.split(".")[-1]
        base_leaf = leaf.split(":")[0]
        shp_want = want[base_leaf] if ":" not in leaf else ((ff, d) if base_leaf == "w_up" else (inner, d))
        if tuple(t.shape) != shp_want:
            raise ValueError(f"checkpoint tensor {hf_name} has shape {tuple(t.shape)}, the config implies {shp_want}")
        W[n] = t.to(device=device, dtype=weight_dtype if _t5.is_matrix(n.split(":")[0]) else torch.float32).contiguous()
        base = n.split(":")[0]
        if base != n:
            blocks = [f"{base}:{i}" for i in range(3 if base.endswith("w_qkv") else 2)]
            if all(b in W for b in blocks):
                W[base] = torch.cat([W.pop(b) for b in blocks], dim=0).contiguous()
    if "lm_head" not in W:      # tied checkpoints store the shared embedding only
        W["lm_head"] = W["wte"]
    expected = [n for n in _t5.slot_names(shape) if n is not None and n not in ("enc_bias", "dec_bias")] + ["enc_rel_bias", "dec_rel_bias"]
    missing = [n for n in expected if n not in W]
    if missing:
        raise ValueError(f"checkpoint is missing {len(missing)} tensors, e.g. {missing[:4]}")
    return W


def iter_checkpoint_tensors(path: str) -> Iterator[Tuple[str, torch.Tensor]]:
    st = sorted(glob.glob(os.path.join(path, "*.safetensors")))
    if st:
        from safetensors import safe_open
        for f in st:
            with safe_open(f, framework="pt", device="cpu") as fh:
                for k in fh.keys():
                    yield k, fh.get_tensor(k)
        return
    bins = sorted(glob.glob(os.path.join(path, "pytorch_model*.bin")))
    if not bins:
        raise FileNotFoundError(f"no *.safetensors or pytorch_model*.bin under {path}")
    for f in bins:
        sd = torch.load(f, map_location="cpu", weights_only=True)
        for k, v in sd.items():
            yield k, v
        del sd


def resolve_local(model_name: str) -> Optional[str]:
    """Directory of a checkpoint for ``model_name``: the name itself if it is a directory, else
    ``$B2S_CHECKPOINT_ROOT/<name>`` (with or without the organisation prefix).  None when there is no local copy."""
    cands = [model_name]
    root = os.environ.get("B2S_CHECKPOINT_ROOT")
    if root:
        cands += [os.path.join(root, model_name), os.path.join(root, model_name.split("/")[-1]),
                  os.path.join(root, model_name.replace("/", "--"))]
    for c in cands:
        if os.path.isfile(os.path.join(c, "config.json")):
            return c
    return None


def load_checkpoint(path: str, device="cuda:0", weight_dtype=torch.float16, n_layer: int = 0):
    """(shape, weights on ``device``) from a local HF checkpoint directory.  ``n_layer`` > 0 keeps only the first
    layers (smoke runs).  A T5 checkpoint comes back as (``t5.EncDecShape``, weights in the layout of ``t5.slot_names``)."""
    with open(os.path.join(path, "config.json")) as f:
        cfg = json.load(f)
    if cfg.get("model_type") in ("t5", "mt5"):
        from dataclasses import replace
        shape = t5_shape_from_hf_config(cfg, name=os.path.basename(os.path.normpath(path)))
        if n_layer:
            shape = replace(shape, n_enc_layer=min(n_layer, shape.n_enc_layer), n_dec_layer=min(n_layer, shape.n_dec_layer))

        def wanted_t5():
            for k, v in iter_checkpoint_tensors(path):
                m = _T5_RE.match(k)
                if m and int(m.group(2)) >= (shape.n_enc_layer if m.group(1) == "encoder" else shape.n_dec_layer):
                    continue
                yield k, v
        return shape, t5_flat_from_hf(shape, wanted_t5(), device=device, weight_dtype=weight_dtype)
    shape = shape_from_hf_config(cfg, name=os.path.basename(os.path.normpath(path)))
    if n_layer:
        shape = shape.with_layers(min(n_layer, shape.n_layer))

    def wanted():
        for k, v in iter_checkpoint_tensors(path):
            m = _LAYER_RE.match(k)
            if m and int(m.group(1)) >= shape.n_layer:
                continue
            yield k, v
    norm_head = cfg.get("model_type") == "baichuan" and int(cfg.get("vocab_size", 0)) > 100000     # Baichuan2 (Baichuan-1: plain head)
    return shape, flat_from_hf(shape, wanted(), device=device, weight_dtype=weight_dtype, norm_head=norm_head)


def load_tokenizer(path: str):
    from transformers import AutoTokenizer
    tok = AutoTokenizer.from_pretrained(path, local_files_only=True)
    if tok.pad_token is None:                       # reference run_base_vs_instruct_100q.py:420-421
        tok.pad_token = tok.eos_token
    return tok
