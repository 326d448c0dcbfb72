"""SURVEY 8f row N2 on the GPU: a checkpoint + tokenizer written to disk by transformers, loaded through the entry
points' load_model, scored through the C-ABI -- against the reference call sequence run on the HF model that
AutoModelForCausalLM.from_pretrained builds from the same directory."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

PROMPTS = ['Question: Is "soup" a "beverage"? Answer either \'Yes\' or \'No\', without any other text.\nAnswer:',
           'Question: Is a "tweet" a "publication"? Answer either \'Yes\' or \'No\'.\nAnswer: Yes.\n\nQuestion: Is a "screenshot" a "photograph"?\nAnswer:',
           'Is the soup a beverage?']


@pytest.mark.parametrize("maker", ["tiny_falcon", "tiny_neox_parallel", "tiny_llama", "tiny_opt", "tiny_bloom"])
def test_local_checkpoint_end_to_end_vs_hf(tmp_path, monkeypatch, maker):
    from transformers import AutoModelForCausalLM, AutoTokenizer
    from llm_interpretation_replication_b200 import scorer as S, shapes, weights
    from llm_interpretation_replication_b200.entrypoints import common
    from llm_interpretation_replication_b200.tokenizer import HashTokenizer
    from oracle import hf_models, scorer_oracle
    from test_checkpoint_cpu import _write_tokenizer

    shp = getattr(shapes, maker)(vocab=64)
    W = weights.init_weights(shp, seed=31)
    d = tmp_path / "someorg" / "some-model"
    d.mkdir(parents=True)
    hf_models.build_hf_model(shp, W).save_pretrained(str(d))
    _write_tokenizer(str(d))
    monkeypatch.setenv("B2S_CHECKPOINT_ROOT", str(tmp_path))

    scorer, tok = common.load_model("someorg/some-model", mode=shapes.MODE_FP32_CHECK, max_tokens=4096, max_seqs=16)
    assert not isinstance(tok, HashTokenizer) and scorer.shape.vocab == 64 and scorer.shape.n_layer == shp.n_layer
    hf = AutoModelForCausalLM.from_pretrained(str(d), torch_dtype=torch.float32).eval()
    hftok = AutoTokenizer.from_pretrained(str(d))
    if hftok.pad_token is None:
        hftok.pad_token = hftok.eos_token
    for variant in ("100q", "instruct"):
        ours = S.score_prompts(scorer, tok, PROMPTS, variant=variant, gen_tokens=50)
        for p, o in zip(PROMPTS, ours):
            ref = scorer_oracle.get_yes_no_logprobs(hf, hftok, p, "cpu", variant)
            assert abs(o["relative_prob"] - ref["relative_prob"]) <= 1e-5, (variant, p)
            assert abs(o["yes_prob"] - ref["yes_prob"]) <= 1e-5 and abs(o["no_prob"] - ref["no_prob"]) <= 1e-5
            assert o["completion"] == ref["completion"], (variant, p)
            if variant == "instruct":
                assert o["position_found"] == ref["position_found"] and o["yes_no_found"] == ref["yes_no_found"]
    common.unload(scorer)

    # a 16-bit run of the same checkpoint stays inside the 2e-3 gate
    scorer, tok = common.load_model("someorg/some-model", mode=shapes.MODE_FP16, max_tokens=4096, max_seqs=16)
    ours = S.score_prompts(scorer, tok, PROMPTS, variant="instruct")
    for p, o in zip(PROMPTS, ours):
        ref = scorer_oracle.get_yes_no_logprobs(hf, hftok, p, "cpu", "instruct")
        if o["position_found"] == ref["position_found"]:
            assert abs(o["relative_prob"] - ref["relative_prob"]) <= 2e-3
    common.unload(scorer)


def test_unsupported_checkpoint_becomes_nan_rows(tmp_path, monkeypatch):
    """A local checkpoint of an architecture outside the supported families behaves like a model that fails to load
    in the reference: NaN rows, the sweep goes on (run_base_vs_instruct_100q.py:493-507)."""
    import json
    from llm_interpretation_replication_b200.entrypoints import run_base_vs_instruct_100q as ep
    d = tmp_path / "org" / "llama"
    d.mkdir(parents=True)
    (d / "config.json").write_text(json.dumps({"model_type": "t5", "hidden_size": 64}))
    monkeypatch.setenv("B2S_CHECKPOINT_ROOT", str(tmp_path))
    rows = ep.run_model_on_prompts("org/llama", ["Is a tweet a publication?"], is_base_model=False)
    assert len(rows) == 1 and rows[0]["success"] is False and np.isnan(rows[0]["relative_prob"])
    assert rows[0]["completion"].startswith("MODEL_ERROR")
