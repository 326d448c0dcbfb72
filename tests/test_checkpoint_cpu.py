"""SURVEY 8f row N2 on the CPU: local HF checkpoints (safetensors / .bin) and tokenizers -> the flat weight layout."""
import json
import os

import numpy as np
import pytest
import torch

from llm_interpretation_replication_b200 import checkpoint, shapes, weights
from oracle import hf_models


def _save_hf(shape, W, path, safe=True):
    model = hf_models.build_hf_model(shape, W)
    model.save_pretrained(path, safe_serialization=safe)
    return model


@pytest.mark.parametrize("maker", ["tiny_falcon", "tiny_neox_parallel", "tiny_neox_sequential", "tiny_llama", "tiny_mistral",
                                   "tiny_opt", "tiny_bloom", "tiny_qwen"])
@pytest.mark.parametrize("safe", [True, False], ids=["safetensors", "bin"])
def test_checkpoint_round_trip(tmp_path, maker, safe):
    """A checkpoint written by transformers itself comes back as the same shape and bit-identical tensors."""
    shp = getattr(shapes, maker)()
    W = weights.init_weights(shp, seed=7)
    _save_hf(shp, W, str(tmp_path), safe=safe)
    shp2, W2 = checkpoint.load_checkpoint(str(tmp_path), device="cpu", weight_dtype=torch.float32)
    for f in ("arch", "n_layer", "d_model", "n_head", "n_kv_head", "head_dim", "d_ff", "vocab", "rotary_dim", "rope_theta",
              "norm_eps", "parallel_residual", "n_ln_per_layer", "has_bias", "tied_head", "qkv_layout", "eos_id", "max_pos",
              "norm_type", "mlp_type", "pos_type", "pos_offset", "act_type", "emb_norm", "qkv_bias"):
        assert getattr(shp2, f) == getattr(shp, f), f
    names = checkpoint.expected_names(shp)
    assert set(names) <= set(W2)
    for n in names:
        assert torch.equal(W2[n], W[n]), n
    if shp.tied_head:
        assert W2["lm_head"] is W2["wte"]
    # truncated load keeps the first layer only
    shp1, W1 = checkpoint.load_checkpoint(str(tmp_path), device="cpu", weight_dtype=torch.bfloat16, n_layer=1)
    assert shp1.n_layer == 1 and "l1.w_qkv" not in W1 and W1["l0.w_qkv"].dtype == torch.bfloat16
    assert W1["l0.ln1_g"].dtype == torch.float32


def test_config_mapping_matches_registry():
    """The public config.json key sets of the sweep's models map onto the registry shapes."""
    old_falcon = dict(model_type="RefinedWebModel", n_layer=32, n_head=71, hidden_size=4544, vocab_size=65024, multi_query=True,
                      parallel_attn=True, bias=False, alibi=False, layer_norm_epsilon=1e-5, eos_token_id=11, bos_token_id=11)
    new_falcon = dict(model_type="falcon", num_hidden_layers=32, num_attention_heads=71, hidden_size=4544, vocab_size=65024,
                      multi_query=True, parallel_attn=True, bias=False, alibi=False, new_decoder_architecture=False,
                      layer_norm_epsilon=1e-5, eos_token_id=11, max_position_embeddings=2048, rope_theta=10000.0)
    stablelm = dict(model_type="gpt_neox", hidden_size=6144, num_hidden_layers=16, num_attention_heads=48, intermediate_size=24576,
                    rotary_pct=0.25, rotary_emb_base=10000, layer_norm_eps=1e-5, use_parallel_residual=True, vocab_size=50432,
                    max_position_embeddings=4096, eos_token_id=0, hidden_act="gelu", tie_word_embeddings=False)
    redpajama = dict(model_type="gpt_neox", hidden_size=2560, num_hidden_layers=32, num_attention_heads=32, intermediate_size=10240,
                     rotary_pct=1.0, rotary_emb_base=10000, layer_norm_eps=1e-5, use_parallel_residual=False, vocab_size=50432,
                     max_position_embeddings=2048, eos_token_id=0, hidden_act="gelu", tie_word_embeddings=False)
    mistral = dict(model_type="mistral", hidden_size=4096, num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=8,
                   intermediate_size=14336, vocab_size=32768, rms_norm_eps=1e-5, rope_theta=1000000.0, sliding_window=None,
                   max_position_embeddings=32768, hidden_act="silu", tie_word_embeddings=False, eos_token_id=2, bos_token_id=1)
    llama2 = dict(model_type="llama", hidden_size=4096, num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=32,
                  intermediate_size=11008, vocab_size=32000, rms_norm_eps=1e-5, max_position_embeddings=4096, hidden_act="silu",
                  tie_word_embeddings=False, eos_token_id=2, bos_token_id=1)
    opt = dict(model_type="opt", hidden_size=2048, num_hidden_layers=24, ffn_dim=8192, num_attention_heads=32, vocab_size=50272,
               max_position_embeddings=2048, do_layer_norm_before=True, word_embed_proj_dim=2048, activation_function="relu",
               enable_bias=True, eos_token_id=2, tie_word_embeddings=True)
    bloom = dict(model_type="bloom", hidden_size=4096, n_layer=30, n_head=32, vocab_size=250880, layer_norm_epsilon=1e-5,
                 apply_residual_connection_post_layernorm=False, eos_token_id=2, seq_length=2048)
    assert checkpoint.shape_from_hf_config(opt, "opt-iml-1.3b") == shapes.opt_iml_1_3b()
    assert checkpoint.shape_from_hf_config(bloom, "bloomz-7b1") == shapes.bloomz_7b1()
    for cfg, ref in ((old_falcon, shapes.falcon_7b()), (new_falcon, shapes.falcon_7b()), (stablelm, shapes.stablelm_alpha_7b()),
                     (redpajama, shapes.redpajama_3b()), (mistral, shapes.mistral_7b()), (llama2, shapes.llama2_7b())):
        got = checkpoint.shape_from_hf_config(cfg, ref.name)
        assert got == ref, (got, ref)
    for bad in (dict(model_type="t5"), dict(model_type="opt", hidden_size=1024, word_embed_proj_dim=512, num_hidden_layers=24,
                                               ffn_dim=4096, num_attention_heads=16, vocab_size=50272, do_layer_norm_before=False),
                dict(model_type="mistral", hidden_size=64, num_attention_heads=2, num_hidden_layers=1,
                                                  intermediate_size=128, vocab_size=32, sliding_window=1024),
                dict(model_type="falcon", hidden_size=8192, num_attention_heads=128, num_hidden_layers=60,
                                                  vocab_size=65024, new_decoder_architecture=True),
                dict(model_type="gpt_neox", hidden_size=64, num_attention_heads=1, num_hidden_layers=1, intermediate_size=256,
                     vocab_size=10, hidden_act="gelu_new")):
        with pytest.raises(KeyError):
            checkpoint.shape_from_hf_config(bad, "x")


def test_wrong_or_missing_tensors_are_rejected(tmp_path):
    shp = shapes.tiny_falcon()
    W = weights.init_weights(shp, seed=1)
    sd = hf_models.to_hf_state_dict(shp, W)
    bad = dict(sd); bad["transformer.h.0.mlp.dense_h_to_4h.weight"] = torch.zeros(3, 3)
    with pytest.raises(ValueError, match="shape"):
        checkpoint.flat_from_hf(shp, bad)
    short = {k: v for k, v in sd.items() if "h.1.self_attention.dense" not in k}
    with pytest.raises(ValueError, match="missing"):
        checkpoint.flat_from_hf(shp, short)
    # buffers HF checkpoints carry but the scorer does not use are ignored
    extra = dict(sd); extra["transformer.h.0.self_attention.rotary_emb.inv_freq"] = torch.zeros(32)
    assert set(checkpoint.flat_from_hf(shp, extra)) == set(checkpoint.flat_from_hf(shp, sd))
    with pytest.raises(FileNotFoundError):
        (tmp_path / "config.json").write_text(json.dumps(hf_models.hf_config(shp).to_dict()))
        checkpoint.load_checkpoint(str(tmp_path), device="cpu")


def _write_tokenizer(path, vocab_size=256):
    """A real (fast) HF tokenizer on disk: word-level over a closed vocabulary, ' Yes' / ' No' as single tokens."""
    from tokenizers import Tokenizer, models, pre_tokenizers
    from transformers import PreTrainedTokenizerFast
    words = ["<eos>", "<unk>", "Yes", "No", "Is", "a", "the", "Question", "Answer", ":", "?", ".", ",", "'", '"', "soup",
             "beverage", "tweet", "publication", "either", "or", "without", "any", "other", "text", "screenshot", "photograph"]
    vocab = {w: i for i, w in enumerate(words)}
    tk = Tokenizer(models.WordLevel(vocab, unk_token="<unk>"))
    tk.pre_tokenizer = pre_tokenizers.Sequence([pre_tokenizers.Whitespace()])
    fast = PreTrainedTokenizerFast(tokenizer_object=tk, eos_token="<eos>", unk_token="<unk>")
    fast.save_pretrained(path)
    return vocab


def test_tokenizer_pass_through_and_candidate_ids(tmp_path):
    from llm_interpretation_replication_b200 import scorer as S
    vocab = _write_tokenizer(str(tmp_path))
    tok = checkpoint.load_tokenizer(str(tmp_path))
    assert tok.pad_token == tok.eos_token                      # reference :420-421
    ids = tok('Is a "tweet" a "publication"?').input_ids
    assert ids[0] == vocab["Is"] and vocab["tweet"] in ids
    for variant in ("100q", "instruct"):
        a, b = S.candidate_ids(tok, variant)
        assert (a, b) == (vocab["Yes"], vocab["No"])
    assert tok.decode([vocab["Yes"], vocab["<eos>"]], skip_special_tokens=True).strip() == "Yes"


def test_resolve_local(tmp_path, monkeypatch):
    d = tmp_path / "tiiuae" / "falcon-7b"
    d.mkdir(parents=True)
    (d / "config.json").write_text("{}")
    assert checkpoint.resolve_local("tiiuae/falcon-7b") is None
    monkeypatch.setenv("B2S_CHECKPOINT_ROOT", str(tmp_path))
    assert checkpoint.resolve_local("tiiuae/falcon-7b") == str(d)
    assert checkpoint.resolve_local(str(d)) == str(d)
    assert checkpoint.resolve_local("tiiuae/falcon-40b") is None



def test_qwen1_and_baichuan2_key_maps(tmp_path):
    """Qwen-7B (QWenLMHeadModel) and Baichuan2-7B (BaichuanForCausalLM) ship as custom code this image cannot instantiate, so
    only the MAPPING is tested here: a state dict laid out under their published key names and config keys loads into the same
    flat weights as the Llama / Qwen2 layout of the same tensors (whose math is pinned against HF), with Qwen's w2 as the gate,
    Baichuan's packed W_pack, and Baichuan2's lm_head rows L2-normalised once at load."""
    import json
    from safetensors.torch import save_file
    # Qwen-1
    shp = shapes.tiny_qwen()
    W = weights.init_weights(shp, seed=21)
    ff = shp.d_ff
    sd = {"transformer.wte.weight": W["wte"], "transformer.ln_f.weight": W["lnf_g"], "lm_head.weight": W["lm_head"]}
    for l in range(shp.n_layer):
        p, h = f"l{l}.", f"transformer.h.{l}."
        sd[h + "ln_1.weight"] = W[p + "ln1_g"]; sd[h + "ln_2.weight"] = W[p + "ln2_g"]
        sd[h + "attn.c_attn.weight"] = W[p + "w_qkv"]; sd[h + "attn.c_attn.bias"] = W[p + "b_qkv"]
        sd[h + "attn.c_proj.weight"] = W[p + "w_dense"]
        sd[h + "mlp.w2.weight"] = W[p + "w_up"][:ff]          # gate: c_proj(w1(x) * silu(w2(x)))
        sd[h + "mlp.w1.weight"] = W[p + "w_up"][ff:]
        sd[h + "mlp.c_proj.weight"] = W[p + "w_down"]
    d = tmp_path / "qwen"
    d.mkdir()
    save_file({k: v.contiguous().clone() for k, v in sd.items()}, str(d / "model.safetensors"))
    json.dump(dict(model_type="qwen", architectures=["QWenLMHeadModel"], hidden_size=shp.d_model, num_hidden_layers=shp.n_layer,
                   num_attention_heads=shp.n_head, kv_channels=shp.head_dim, intermediate_size=2 * ff, vocab_size=shp.vocab,
                   layer_norm_epsilon=1e-6, rotary_emb_base=10000, rotary_pct=1.0, seq_length=2048, no_bias=True,
                   use_dynamic_ntk=True, use_logn_attn=True, tie_word_embeddings=False, eos_token_id=2), open(d / "config.json", "w"))
    shp2, W2 = checkpoint.load_checkpoint(str(d), device="cpu", weight_dtype=torch.float32)
    assert (shp2.d_ff, shp2.qkv_bias, shp2.norm_type, shp2.mlp_type, shp2.n_kv_head, shp2.head_dim) == (ff, True, shp.norm_type, shp.mlp_type, shp.n_head, 64)
    assert set(W2) == set(W) and all(torch.equal(W[k], W2[k]) for k in W)
    # Baichuan2
    shp = shapes.tiny_llama(n_kv_head=4)
    W = weights.init_weights(shp, seed=22)
    ff, V = shp.d_ff, 100352
    big = torch.randn(V, shp.d_model, generator=torch.Generator().manual_seed(1)) * 3.0
    sd = {"model.embed_tokens.weight": big, "model.norm.weight": W["lnf_g"], "lm_head.weight": big * 0.5 + 0.1}
    for l in range(shp.n_layer):
        p, h = f"l{l}.", f"model.layers.{l}."
        sd[h + "input_layernorm.weight"] = W[p + "ln1_g"]; sd[h + "post_attention_layernorm.weight"] = W[p + "ln2_g"]
        sd[h + "self_attn.W_pack.weight"] = W[p + "w_qkv"]; sd[h + "self_attn.o_proj.weight"] = W[p + "w_dense"]
        sd[h + "mlp.gate_proj.weight"] = W[p + "w_up"][:ff]; sd[h + "mlp.up_proj.weight"] = W[p + "w_up"][ff:]
        sd[h + "mlp.down_proj.weight"] = W[p + "w_down"]
    d = tmp_path / "baichuan"
    d.mkdir()
    save_file({k: v.contiguous().clone() for k, v in sd.items()}, str(d / "model.safetensors"))
    json.dump(dict(model_type="baichuan", architectures=["BaichuanForCausalLM"], hidden_size=shp.d_model, intermediate_size=ff,
                   num_attention_heads=shp.n_head, num_hidden_layers=shp.n_layer, rms_norm_eps=1e-6, vocab_size=V,
                   model_max_length=4096, tie_word_embeddings=False, bos_token_id=1, eos_token_id=2, pad_token_id=0),
              open(d / "config.json", "w"))
    shp2, W2 = checkpoint.load_checkpoint(str(d), device="cpu", weight_dtype=torch.float32)
    assert shp2.vocab == V and not shp2.qkv_bias and shp2.n_kv_head == shp.n_head and shp2.max_pos == 4096
    assert torch.equal(W2["l1.w_qkv"], W["l1.w_qkv"]) and torch.equal(W2["l0.w_up"], W["l0.w_up"])
    assert torch.allclose(W2["lm_head"].norm(dim=-1), torch.ones(V), atol=1e-5)                     # NormHead
    assert torch.allclose(W2["lm_head"], torch.nn.functional.normalize(sd["lm_head.weight"], dim=-1))
    with pytest.raises(KeyError):
        checkpoint.shape_from_hf_config(dict(model_type="baichuan", hidden_size=5120, num_hidden_layers=40, num_attention_heads=40,
                                             intermediate_size=13696, vocab_size=125696), "13b")
