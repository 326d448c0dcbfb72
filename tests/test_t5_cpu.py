"""Encoder-decoder (T5) branch on CPU: the oracle's restatement against the HF module the reference loads, the oracle against
the goldens produced by the reference's own `is_encoder_decoder` branch, and the host logic (bias tables, slot layout)."""
import json
import os

import pytest
import torch

from conftest import GOLDEN


def _cases():
    return json.load(open(os.path.join(GOLDEN, "t5_golden.json")))["cases"]


def _tok(shp, row=None):
    from llm_interpretation_replication_b200.tokenizer import HashTokenizer
    tok = HashTokenizer(shp.vocab, eos_token_id=shp.eos_id, append_eos=True, pad_token_id=shp.pad_id)
    if row is not None and row.get("cand_override") is not None:
        tok.overrides = {"Yes": row["cand_override"][0], "No": row["cand_override"][1]}
    return tok


@pytest.mark.parametrize("name", ["tiny-t5", "tiny-t5-relu"])
def test_t5_restatement_matches_hf_generate(name):
    from llm_interpretation_replication_b200 import t5
    from oracle import hf_models, t5_ref
    shp = t5.REGISTRY[name]()
    W = t5.init_weights(shp, seed=7)
    model = hf_models.build_hf_t5(shp, W)
    g = torch.Generator().manual_seed(3)
    for L in (4, 37, 150):     # 150: relative distances beyond relative_attention_max_distance (128)
        ids = torch.randint(2, shp.vocab, (1, L), generator=g)
        ids[0, -1] = shp.eos_id
        with torch.no_grad():
            out = model.generate(ids, max_new_tokens=10, output_scores=True, return_dict_in_generate=True, do_sample=False,
                                 pad_token_id=shp.pad_id, eos_token_id=shp.eos_id)
        rows, gen = t5_ref.greedy_scores(shp, W, ids[0].tolist(), 10, shp.eos_id)
        assert out.sequences[0].tolist()[1:] == gen and len(out.scores) == len(rows)
        for a, b in zip(out.scores, rows):
            assert (a[0] - b).abs().max().item() < 2e-5


def test_t5_bias_table_matches_hf_compute_bias():
    from llm_interpretation_replication_b200 import t5
    from oracle import hf_models
    shp = t5.tiny_t5()
    W = t5.init_weights(shp, seed=11)
    model = hf_models.build_hf_t5(shp, W)
    R = shp.max_pos - 1
    for stack, key, bidir in ((model.encoder, "enc_rel_bias", True), (model.decoder, "dec_rel_bias", False)):
        att = stack.block[0].layer[0].SelfAttention
        L = 200
        with torch.no_grad():
            ref = att.compute_bias(L, L)[0]                      # [nh, q, k]
        tab = t5.bias_table(W[key], bidir, shp)                  # [nh, 2R+1] by (k - q)
        q = torch.arange(L)[:, None]
        k = torch.arange(L)[None, :]
        assert torch.equal(tab[:, (k - q) + R], ref)


def test_t5_slot_layout_matches_header():
    import re
    from llm_interpretation_replication_b200 import t5
    hdr = open(os.path.join(os.path.dirname(GOLDEN), "..", "include", "b200score.h")).read()
    n = {k: int(v) for k, v in re.findall(r"(B2S_T5_NGLOBAL|B2S_T5_ENC_PER_LAYER|B2S_T5_DEC_PER_LAYER) = (\d+)", hdr)}
    assert (n["B2S_T5_NGLOBAL"], n["B2S_T5_ENC_PER_LAYER"], n["B2S_T5_DEC_PER_LAYER"]) == (t5.N_GLOBAL, t5.N_PER_ENC, t5.N_PER_DEC)
    shp = t5.tiny_t5_relu()
    names = t5.slot_names(shp)
    assert len(names) == t5.N_GLOBAL + t5.N_PER_ENC * shp.n_enc_layer + t5.N_PER_DEC * shp.n_dec_layer
    enc = re.search(r"enum \{ B2S_T5E_LN1_G = 0, ([^}]*)\}", hdr).group(1)
    dec = re.search(r"enum \{ B2S_T5D_LN1_G = 0, ([^}]*)\}", hdr).group(1)
    assert ["ln1_g"] + [s.strip()[len("B2S_T5E_"):].lower() for s in enc.split(",") if s.strip()] == [s for s in t5.ENC_SLOTS if s]
    assert ["ln1_g"] + [s.strip()[len("B2S_T5D_"):].lower() for s in dec.split(",") if s.strip()] == [s for s in t5.DEC_SLOTS if s]
    W = t5.init_weights(shp, seed=1)
    have = {k for k in W} - {"enc_rel_bias", "dec_rel_bias"} | {"enc_bias", "dec_bias"}
    assert have == {s for s in names if s}


@pytest.mark.parametrize("case_idx", [0, 1])
def test_t5_oracle_matches_reference_golden(case_idx):
    """The restated scan + forward reproduce what the reference's own functions returned (top-5 script, top-2 script)."""
    from llm_interpretation_replication_b200 import t5
    from llm_interpretation_replication_b200.scorer import candidate_ids
    from oracle import t5_ref
    case = _cases()[case_idx]
    shp = t5.REGISTRY[case["shape"]]()
    W = t5.init_weights(shp, seed=case["seed"])
    hits = 0
    for row in case["rows"][::2]:
        tok = _tok(shp, row)
        ids = tok(row["prompt"]).input_ids
        assert ids[-1] == shp.eos_id and len(ids) == row["n_tokens"]
        a, b = candidate_ids(tok, "instruct", "Yes", "No")
        o5 = t5_ref.score_ids(shp, W, ids, a, b, topk=5)
        o2 = t5_ref.score_ids(shp, W, ids, a, b, topk=2)
        g5, g2, gb = row["ref_100q"], row["ref_instruct"], row["ref_base_vs"]
        assert abs(o5["p_a"] - g5["yes_prob"]) < 1e-6 and abs(o5["p_b"] - g5["no_prob"]) < 1e-6
        assert abs(o5["rel"] - g5["relative_prob"]) < 1e-5
        assert o2["pos_found"] == g2["position_found"] == gb["position_found"] and o2["found"] == g2["yes_no_found"]
        assert abs(o2["rel"] - g2["relative_prob"]) < 1e-5
        assert abs(o2["p_a"] - gb["yes_prob"]) < 1e-6
        hits += o2["found"]
    assert hits >= 3


def test_t5_tied_scaled_head_restatement_and_checkpoint_round_trip(tmp_path):
    """Original-T5 head (tied + d^-0.5 scaling) in the oracle against HF, and HF save_pretrained -> checkpoint.load_checkpoint
    gives back the flat weights and the shape (the scaling flag is read from either generation of config keys)."""
    from dataclasses import replace
    from llm_interpretation_replication_b200 import t5, checkpoint
    from oracle import hf_models, t5_ref
    for shp in (replace(t5.tiny_t5_relu(), tied_head=True), t5.tiny_t5()):
        W = t5.init_weights(shp, seed=13)
        model = hf_models.build_hf_t5(shp, W)
        ids = torch.randint(2, shp.vocab, (1, 30), generator=torch.Generator().manual_seed(1))
        with torch.no_grad():
            out = model.generate(ids, max_new_tokens=4, output_scores=True, return_dict_in_generate=True, do_sample=False,
                                 pad_token_id=shp.pad_id, eos_token_id=shp.eos_id)
        rows, gen = t5_ref.greedy_scores(shp, W, ids[0].tolist(), 4, shp.eos_id)
        assert max((a[0] - b).abs().max().item() for a, b in zip(out.scores, rows)) < 2e-5
        d = str(tmp_path / shp.name) + ("-tied" if shp.tied_head else "")
        model.save_pretrained(d)
        shp2, W2 = checkpoint.load_checkpoint(d, device="cpu", weight_dtype=torch.float32)
        assert replace(shp2, name=shp.name, max_pos=shp.max_pos) == shp
        assert set(W2) == set(W) and all(torch.equal(W[k], W2[k]) for k in W)
        shp1, W1 = checkpoint.load_checkpoint(d, device="cpu", weight_dtype=torch.float32, n_layer=1)
        assert (shp1.n_enc_layer, shp1.n_dec_layer) == (1, 1) and "e1.w_qkv" not in W1 and "d0.w_ckv" in W1
    hub_v11 = dict(model_type="t5", d_model=768, d_kv=64, d_ff=2048, num_layers=12, num_decoder_layers=12, num_heads=12,
                   vocab_size=32128, feed_forward_proj="gated-gelu", tie_word_embeddings=False, layer_norm_epsilon=1e-6,
                   relative_attention_num_buckets=32, eos_token_id=1, pad_token_id=0, decoder_start_token_id=0)
    assert replace(checkpoint.t5_shape_from_hf_config(hub_v11, "x"), name="t5-v1_1-base") == t5.t5_v1_1_base()
    assert checkpoint.t5_shape_from_hf_config(dict(hub_v11, tie_word_embeddings=True, feed_forward_proj="relu"), "x").tied_head
    with pytest.raises(KeyError):
        checkpoint.t5_shape_from_hf_config(dict(hub_v11, d_kv=128), "x")
