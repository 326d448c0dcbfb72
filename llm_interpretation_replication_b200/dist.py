"""Prompt sharding across ranks + the single result gather.

Every prompt is an independent sequence and the weights are replicated, so the path shards by prompt with
no data-path collective; the only exchange is one all-gather of the per-prompt result rows at the end
(NCCL over NVLink on the GPU box; the same code runs over gloo in the CPU tests).
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np
import torch


def shard_prompts(lengths: Sequence[int], world_size: int) -> List[List[int]]:
    """Sort by length (longest first) and deal in a serpentine over the ranks, so that every rank gets the
    same number of prompts (+-1) and nearly the same token count in every length range."""
    order = sorted(range(len(lengths)), key=lambda i: (-lengths[i], i))
    shards: List[List[int]] = [[] for _ in range(world_size)]
    for k, i in enumerate(order):
        rnd, pos = divmod(k, world_size)
        r = pos if rnd % 2 == 0 else world_size - 1 - pos
        shards[r].append(i)
    return shards


def gather_rows(local_rows: torch.Tensor, local_index: Sequence[int], n_total: int, world_size: int,
                group=None) -> torch.Tensor:
    """All-gather [n_local, C] result rows from every rank and scatter them back into prompt order.
    ``local_rows`` lives on the device the process group communicates on (cuda for nccl, cpu for gloo)."""
    import torch.distributed as dist
    C = local_rows.shape[1]
    per = (n_total + world_size - 1) // world_size
    dev = local_rows.device
    buf = torch.full((per, C + 1), -1.0, dtype=torch.float32, device=dev)
    n = local_rows.shape[0]
    buf[:n, :C] = local_rows
    buf[:n, C] = torch.as_tensor(np.asarray(local_index, dtype=np.float32), device=dev)   # exact for < 2^24 prompts
    out = torch.empty((world_size * per, C + 1), dtype=torch.float32, device=dev)
    dist.all_gather_into_tensor(out, buf, group=group)
    idx = out[:, C].long()
    keep = idx >= 0
    result = torch.empty((n_total, C), dtype=torch.float32, device=dev)
    result[idx[keep]] = out[keep, :C]
    return result


def score_ids_sharded(scorer, id_lists, cand, topk, max_lookahead, eos_id, group=None, gen_tokens=None, nonfinite="nan"):
    """Each rank scores its shard of already-tokenised prompts on its own GPU; one all-gather returns every
    prompt's row (and greedy tokens) to every rank, in the callers' order."""
    import torch.distributed as dist
    g = None if group is True else group
    world = dist.get_world_size(g)
    rank = dist.get_rank(g)
    lens = [len(x) for x in id_lists]
    mine = shard_prompts(lens, world)[rank]
    rows, gens = scorer.score_ids([id_lists[i] for i in mine], cand[mine], topk=topk, max_lookahead=max_lookahead,
                                  eos_id=eos_id, gen_tokens=gen_tokens, nonfinite=nonfinite)
    dev = scorer.device
    local = torch.cat([torch.from_numpy(rows), torch.from_numpy(gens.astype(np.float32))], dim=1).to(dev)
    full = gather_rows(local, mine, len(id_lists), world, group=g).cpu().numpy()
    return full[:, :8].astype(np.float32), full[:, 8:].astype(np.int32)
