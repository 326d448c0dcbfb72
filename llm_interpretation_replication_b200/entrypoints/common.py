"""Shared pieces of the kept entry points: model construction (random-init, no hub access offline),
prompt formats, logging."""
from __future__ import annotations

import json
import os
import sys

import torch

from .. import shapes, t5, weights
from ..scorer import Scorer
from ..tokenizer import HashTokenizer

# The reference always generates max_new_tokens=50 (run_base_vs_instruct_100q.py:338-346, compare_instruct_models.py:
# 274) and decodes them into the `completion` / `model_output` columns; only the first 10 positions are scanned.
COMPLETION_TOKENS = int(os.environ.get("B2S_COMPLETION_TOKENS", "50"))

DATA_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data")

# reference analysis/run_base_vs_instruct_100q.py:455-461
FEW_SHOT = '''Question: Is "soup" a "beverage"? Answer either 'Yes' or 'No', without any other text.
Answer: No.

Question: Is a "tweet" a "publication"? Answer either 'Yes' or 'No', without any other text.
Answer: Yes.

'''


def log_print(*args, **kwargs):
    print(*args, **kwargs)
    sys.stdout.flush()


def survey_questions():
    """The 100 ordinary-meaning questions (hard-coded list of the reference script, :122-230)."""
    d = json.load(open(os.path.join(DATA_DIR, "survey_questions.json")))
    return d["survey1"], d["survey2"]


def format_base(prompt: str) -> str:      # run_base_vs_instruct_100q.py:466-467
    return f"{FEW_SHOT}Question: {prompt} Answer either 'Yes' or 'No', without any other text.\nAnswer:"


def format_instruct(prompt: str) -> str:  # run_base_vs_instruct_100q.py:468-469; compare_instruct_models.py:488
    return f"{prompt} Answer either 'Yes' or 'No', without any other text."


def format_baichuan(prompt: str) -> str:  # compare_instruct_models.py:491-492
    return f"<human>: {prompt} Answer either 'Yes' or 'No', without any other text.\n<bot>:"


def format_perturbation(rephrased_main: str, response_format: str) -> str:   # perturb_prompts.py:215
    return f"{rephrased_main} {response_format}"


def model_seed(model_name: str) -> int:
    import zlib
    return zlib.crc32(model_name.encode()) & 0x7FFFFFFF


def default_mode(model_name: str, local=None) -> int:
    """The fastest compute mode that holds the 2e-3 gate for the model's family at full depth (DESIGN.md 5,
    profiles/r2_parity_by_family_full_depth.jsonl): FP16 for Falcon / GPT-NeoX / OPT / BLOOM; the Llama block (Llama-2,
    Mistral, Qwen, Baichuan2) is an order of magnitude more sensitive to the rounding of q/k/v and gets fp32 attention under
    hi + lo bf16 GEMM inputs; T5 gets fp16 with the split feed-forward activation.  ``B2S_USE_8BIT=1`` = the reference's
    USE_8BIT = True (run_base_vs_instruct_100q.py:84, compare_instruct_models.py:412): LLM.int8 for the decoder-only models
    (emulated: shapes.MODE_INT8_EMU); the encoder-decoder scorer has no such mode and keeps its default."""
    if local is not None:
        with open(os.path.join(local, "config.json")) as f:
            mt = json.load(f).get("model_type", "")
        family = "t5" if mt in ("t5", "mt5") else "llama" if mt in ("llama", "mistral", "qwen", "qwen2", "baichuan") else "other"
    elif model_name in t5.HUB_TO_SHAPE or model_name in t5.REGISTRY:
        family = "t5"
    else:
        mk = shapes.HUB_TO_SHAPE.get(model_name) or shapes.REGISTRY.get(model_name)
        family = "llama" if (mk is not None and mk().arch == shapes.ARCH_LLAMA) else "other"
    if os.environ.get("B2S_USE_8BIT", "0") not in ("", "0") and family != "t5":
        return shapes.MODE_INT8_EMU
    return {"t5": shapes.MODE_FP16_FFN_SPLIT, "llama": shapes.MODE_BF16_SPLIT_ATTN32}.get(family, shapes.MODE_FP16)


def load_model(model_name: str, device=None, mode=None, n_layer: int = 0, max_tokens=None,
               max_seqs: int = 1024):
    """AutoTokenizer/AutoModelForCausalLM.from_pretrained of the reference (:416-450).  A checkpoint that is on disk
    (``model_name`` is a directory, or ``$B2S_CHECKPOINT_ROOT/<model_name>`` exists; see ``checkpoint.py``) is loaded
    with its own tokenizer.  Otherwise (no hub access here) the stand-in is the shape of the named hub model with
    random-init weights seeded by the name and the offline tokenizer.  Raises KeyError for architectures outside the
    supported families, which the callers turn into NaN rows as the reference does for models that fail to load.
    ``max_tokens`` None = the batch is sized from the device memory that is free once the weights are resident
    (``scorer.fit_batch``: a multi-head 7B model needs 0.5 MB of KV cache per prompt token, Falcon 8 KB).
    ``mode`` None = the fastest mode that holds the 2e-3 gate for the family at full depth: FP16 (Falcon, GPT-NeoX, OPT, BLOOM),
    BF16_SPLIT_ATTN32 (the Llama block: Llama-2, Mistral, Qwen, Baichuan2), FP16_FFN_SPLIT (T5, whose feed-forward activations are
    the known case of values beyond fp16's range; see ``t5.EncDecScorer``)."""
    from .. import checkpoint
    from ..scorer import MAX_LOOK_AHEAD
    device = torch.device(device if device is not None else "cuda:0")
    local = checkpoint.resolve_local(model_name)
    if mode is None:
        mode = default_mode(model_name, local)
    # (FP16_FFN_SPLIT wants fp16 matrices except wo in bf16: the checkpoint is read in fp32 and the scorer converts each matrix once)
    dt = {shapes.MODE_FP32_CHECK: torch.float32, shapes.MODE_FP16: torch.float16, shapes.MODE_INT8_EMU: torch.float16,
          shapes.MODE_FP16_FFN_SPLIT: torch.float32}.get(mode, torch.bfloat16)
    if local is not None:
        shp, W = checkpoint.load_checkpoint(local, device=device, weight_dtype=dt, n_layer=n_layer)
        try:
            tok = checkpoint.load_tokenizer(local)
        except Exception as e:   # weights without tokenizer files: offline tokenizer, said out loud
            log_print(f"  no usable tokenizer under {local} ({type(e).__name__}); using the offline HashTokenizer")
            tok = (HashTokenizer(shp.vocab, eos_token_id=shp.eos_id, append_eos=True, pad_token_id=shp.pad_id)
                   if getattr(shp, "is_encoder_decoder", False) else HashTokenizer(shp.vocab, eos_token_id=shp.eos_id).apply_pad_default())
    elif model_name in t5.HUB_TO_SHAPE or model_name in t5.REGISTRY:
        # encoder-decoder stand-in: the named T5 shape, random-init, offline tokenizer in its T5 form (</s> appended, pad 0)
        from dataclasses import replace
        shp = (t5.HUB_TO_SHAPE.get(model_name) or t5.REGISTRY[model_name])()
        if n_layer:
            shp = replace(shp, n_enc_layer=min(n_layer, shp.n_enc_layer), n_dec_layer=min(n_layer, shp.n_dec_layer))
        W = t5.init_weights(shp, seed=model_seed(model_name), device=device, weight_dtype=dt)
        tok = HashTokenizer(shp.vocab, eos_token_id=shp.eos_id, append_eos=True, pad_token_id=shp.pad_id)
    else:
        if model_name in shapes.HUB_TO_SHAPE:
            shp = shapes.HUB_TO_SHAPE[model_name]()
        elif model_name in shapes.REGISTRY:
            shp = shapes.REGISTRY[model_name]()
        else:
            raise KeyError(f"architecture of {model_name} is not supported by the B200 scorer")
        if n_layer:
            shp = shp.with_layers(n_layer)
        W = weights.init_weights_device(shp, seed=model_seed(model_name), device=device, weight_dtype=dt)
        tok = HashTokenizer(shp.vocab, eos_token_id=shp.eos_id).apply_pad_default()
    if getattr(shp, "is_encoder_decoder", False):
        # the cross-attention K/V of every decoder layer stay resident per call: the token budget is capped (t5.EncDecScorer)
        scorer = t5.EncDecScorer(shp, W, mode=mode, device=device, max_tokens=max_tokens, max_seqs=max_seqs,
                                 max_lookahead=max(MAX_LOOK_AHEAD, COMPLETION_TOKENS))
        return scorer, tok
    scorer = Scorer(shp, W, mode=mode, device=device, max_tokens=max_tokens, max_seqs=max_seqs,
                    max_lookahead=max(MAX_LOOK_AHEAD, COMPLETION_TOKENS))
    return scorer, tok


def unload(scorer):
    scorer.close()
    del scorer
    import gc
    gc.collect()
    torch.cuda.empty_cache()


def maybe_init_dist():
    """Under torchrun (WORLD_SIZE > 1): one process per GPU, NCCL process group; returns (rank, world, shard_group)
    where shard_group is what score_prompts(..., shard_group=) takes (None when single-process)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return 0, 1, None
    import torch.distributed as dist
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    return dist.get_rank(), world, True
