"""CPU only (oracle arithmetic, no GPU): how much fp16 / hi+lo-bf16 operand rounding moves relative_prob in the torch restatement
of each family's forward, at the real widths, 2 layers, random-init -- i.e. what ANY correct 16-bit implementation has to
expect.  The Llama block turns out ~3x more sensitive than Falcon's or GPT-NeoX's, and for it rounding q/k/v to fp16 alone
costs as much as rounding every GEMM input (which is why BF16_SPLIT, whose attention is fp16, does not help it and
BF16_SPLIT_ATTN32 does).  Output recorded in profiles/r2_rounding_sensitivity_cpu.txt."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from llm_interpretation_replication_b200 import shapes, weights  # noqa: E402
from oracle import decoder_ref  # noqa: E402

torch.set_num_threads(16)
f16 = lambda t: t.to(torch.float16).to(t.dtype)
split = lambda t: decoder_ref.bf16_round(t) + decoder_ref.bf16_round(t - decoder_ref.bf16_round(t))
ident = lambda t: t


def run(shp, label, n=16):
    W = weights.init_weights(shp, seed=0, randomize_affine=True)
    g = torch.Generator().manual_seed(1)
    res = {"fp16 everywhere": [], "split GEMM inputs + fp16 q/k/v": [], "split GEMM inputs + fp32 q/k/v": [], "fp16 q/k/v only": []}
    for _ in range(n):
        ids = torch.randint(32, shp.vocab - 32, (int(torch.randint(66, 190, (1,), generator=g)),), generator=g)
        rel = lambda lo: (torch.softmax(lo, -1)[1001] / (torch.softmax(lo, -1)[1001] + torch.softmax(lo, -1)[1002])).item()
        r32 = rel(decoder_ref.last_logits(shp, W, ids))
        res["fp16 everywhere"].append(abs(r32 - rel(decoder_ref.last_logits(shp, W, ids, round_act=f16))))
        res["split GEMM inputs + fp16 q/k/v"].append(abs(r32 - rel(decoder_ref.last_logits(shp, W, ids, round_act=split, round_attn=f16))))
        res["split GEMM inputs + fp32 q/k/v"].append(abs(r32 - rel(decoder_ref.last_logits(shp, W, ids, round_act=split, round_attn=ident))))
        res["fp16 q/k/v only"].append(abs(r32 - rel(decoder_ref.last_logits(shp, W, ids, round_act=None, round_attn=f16))))
    for k, v in res.items():
        print(f"{label:32s} {k:34s} mean {np.mean(v):.2e}  max {np.max(v):.2e}", flush=True)


run(shapes.llama2_7b().with_layers(2), "llama-2-7b widths, 2 layers")
run(shapes.stablelm_alpha_7b().with_layers(2), "stablelm-alpha-7b widths, 2 layers")
run(shapes.falcon_7b().with_layers(2), "falcon-7b widths, 2 layers")
