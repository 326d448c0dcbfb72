#!/usr/bin/env python
"""bench.py -- scored prompts/s for the BASELINE.json headline workload.

Workload (BASELINE.json configs[2], named in ``config.workload``): Falcon-7B shape (random-init bf16
weights), 10,000 perturbation-shaped prompts (synthetic token ids with the length histogram SURVEY.md 8d
measured on data/perturbations.json: 66..190 tokens, mean ~111.5), reference look-ahead semantics
(top-k 5, 10 positions).  A "step" is one pass of the hot path over all 10,000 prompts; with N GPUs the
prompts are sharded by rank (total work fixed -> "strong") and the result rows are all-gathered over NCCL.

  python bench.py [--gpus N] [--steps K] [--warmup W]            # our arm (one JSON line on rank 0)
  python bench.py --impl reference [--steps K] [--warmup W]      # the reference's CPU path (oracle port)

Operand format: the default ``--mode fp16`` runs the tensor cores on fp16 weights/activations (fp32 accumulate): it
is the mode that holds the |delta rel_prob| <= 2e-3 per-prompt gate at 32 layers (profiles/r1_full_size_parity.txt:
max 5e-4 vs 2.7e-3 and 79 % pass for ``--mode bf16``), costs ~1 % throughput, and is also the reference's own
non-quantised dtype (run_base_vs_instruct_100q.py:449 torch_dtype=float16).

``value``  : prompts/s with the packed inputs already resident in HBM (b2s_score on device pointers).
``e2e``    : prompts/s through the host-buffer C-ABI call b2s_score_host (pinned host -> device copies of
             every batch and device -> host copies of the result rows inside the timed region).
``roofline``: tensor-core GEMM kernel (dominant): algorithmic 2*M*N*K FLOPs of every launch in the timed
             region / the CUDA-event time of those launches, against the measured sustained bf16 peak.
``cpu_baseline``: the oracle (HF transformers fp32 on the host cores, ``generate(max_new_tokens=10,
             output_scores=True)`` + look-ahead scan = the same ten positions the GPU arm computes) on a bounded,
             length-stratified sample, extrapolated to the 10,000 prompts by the size of each length range.
``other_modes`` / ``complete50`` (N = 1): the same step in the BF16_FAST and BF16_SPLIT operand modes, and the call the
             kept entry points make (b2s_score_complete_host with the reference's 50 generated tokens).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np
import torch

METRIC = "scored prompts/sec (Falcon-7B, 10k perturbations)"
UNIT = "prompts/s"
N_PROMPTS = 10000
TOPK = 5
LOOKAHEAD = 10


# ---- synthetic workload ---------------------------------------------------------------------------
def synthetic_lengths(n=N_PROMPTS, seed=0):
    """Prompt lengths with the histogram of the 10,000 perturbation prompts under the len//4 tokenizer
    (SURVEY.md 8d): ranges [64,96)/[96,128)/[128,160)/[160,192) hold 2664/5257/1726/353 prompts."""
    rs = np.random.RandomState(seed)
    parts = []
    for (lo, hi, cnt) in ((66, 96, 2664), (96, 128, 5257), (128, 160, 1726), (160, 191, 353)):
        parts.append(rs.randint(lo, hi, size=cnt * n // N_PROMPTS))
    L = np.concatenate(parts)
    if len(L) < n:
        L = np.concatenate([L, rs.randint(96, 128, size=n - len(L))])
    rs.shuffle(L)
    return L.astype(np.int64)


def synthetic_prompts(shape, n=N_PROMPTS, seed=0):
    L = synthetic_lengths(n, seed)
    rs = np.random.RandomState(seed + 1)
    ids = [rs.randint(32, shape.vocab - 32, size=int(l)).astype(np.int32) for l in L]
    # five scenarios x 2000 rephrasings, one candidate pair per scenario (perturbations.json target_tokens)
    pairs = np.array([[1001, 1002], [2001, 2002], [3001, 3002], [4001, 4002], [1001, 1002]], dtype=np.int32)
    cand = pairs[(np.arange(n) * 5) // n]
    return ids, cand, L


# ---- clocks ---------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.path = f"/tmp/b2s_clocks_{os.getpid()}.csv"
        self.proc = None
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        for line in open(self.path):
            p = [x.strip() for x in line.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1])); mx.append(float(p[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            out = {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}
        try:
            os.remove(self.path)
        except OSError:
            pass
        return out


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
    return 1400.0, "fallback (B200_PROFILING.md: ~1.4 PFLOP/s sustained)"


# ---- reference arm / cpu baseline (the ONLY code here that touches oracle/) -----------------------------
LENGTH_RANGES = ((66, 96, 2664), (96, 128, 5257), (128, 160, 1726), (160, 191, 353))   # [lo, hi) and prompts of the 10,000 in it

WORKLOAD = "Falcon-7B (random-init) x 10,000 perturbation-shaped prompts, top-5 / 10-position look-ahead (BASELINE configs[2])"


def workload_config(shp, L):
    """The keys that define the workload; identical in both arms."""
    return {"workload": WORKLOAD, "n_prompts": int(len(L)), "total_prompt_tokens": int(L.sum()), "n_layer": shp.n_layer,
            "max_lookahead": LOOKAHEAD, "topk": TOPK,
            "generated_positions": f"{LOOKAHEAD} in both arms: the scanned positions (the reference script goes on to 50 tokens, "
                                   "which only fill its completion column; timed separately as complete50 / full_reference_call_50)"}


def cpu_reference_run(steps, warmup, with_full_call=False):
    """Times the oracle = the reference's get_yes_no_logprobs restated over HF transformers (fp32, all host cores) on
    the SAME work the GPU arm does: prefill + greedy look-ahead over the 10 scanned positions (max_new_tokens=10), top-5
    scan.  One step = one length-stratified group, one prompt from each of the four perturbation length ranges; the
    per-range mean times are extrapolated to the 10,000-prompt set by the number of prompts in each range.  Also timed:
    a prefill-only forward (what a position-0-only scorer would cost) and, with ``with_full_call``, the reference's
    literal call with max_new_tokens=50."""
    from llm_interpretation_replication_b200 import shapes, weights
    from llm_interpretation_replication_b200.tokenizer import HashTokenizer
    from oracle import scorer_oracle, hf_models
    import transformers
    transformers.logging.set_verbosity_error()
    shp = shapes.falcon_7b()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    t0 = time.time()
    W = weights.init_weights(shp, seed=0, randomize_affine=False, bf16_representable=False)
    model = hf_models.build_hf_model(shp, W)
    build_s = time.time() - t0
    tok = HashTokenizer(shp.vocab, eos_token_id=shp.eos_id).apply_pad_default()
    rs = np.random.RandomState(5)
    letters = np.array(list("abcdefghijklmnopqrstuvwxyz ABCDEFGHIJKLMNOPQRSTUVWXYZ,."))

    def prompt_of(n_tokens):
        p = "".join(rs.choice(letters, size=4 * n_tokens))           # 4 bytes per token under the hash tokenizer
        assert len(tok(p).input_ids) == n_tokens
        return p

    per_range = [[] for _ in LENGTH_RANGES]
    tokens_timed = 0
    prefill = []
    for s in range(warmup + steps):
        for r, (lo, hi, _) in enumerate(LENGTH_RANGES):
            n_tok = int(rs.randint(lo, hi))
            prompt = prompt_of(n_tok)
            t1 = time.time()
            res = scorer_oracle.get_yes_no_logprobs(model, tok, prompt, variant="100q", max_new_tokens=LOOKAHEAD)
            dt = time.time() - t1
            assert res["success"]
            if s >= warmup:
                per_range[r].append(dt)
                tokens_timed += n_tok
        if s >= warmup and len(prefill) < 2:                              # prefill-only figure, 111-token prompt
            ids = tok(prompt_of(111), return_tensors="pt").input_ids
            t1 = time.time()
            with torch.no_grad():
                model(ids, logits_to_keep=1)
            prefill.append(time.time() - t1)
    means = [sum(t) / len(t) for t in per_range]
    total_est = sum(m * cnt for m, (_, _, cnt) in zip(means, LENGTH_RANGES))     # seconds for all 10,000 prompts
    n_timed = sum(len(t) for t in per_range)
    out = dict(value=N_PROMPTS / total_est, seconds_per_prompt=total_est / N_PROMPTS, cores=cores, build_s=build_s,
               per_range_s=[round(m, 3) for m in means], prefill_only_s=round(sum(prefill) / len(prefill), 3) if prefill else None,
               step_s=sum(sum(t) for t in per_range) / max(1, steps),
               sample=f"{n_timed} prompts ({n_timed // 4} from each length range [66,96)/[96,128)/[128,160)/[160,191), {tokens_timed} tokens), "
                      f"get_yes_no_logprobs restated over HF transformers {transformers.__version__} fp32 with "
                      f"generate(max_new_tokens={LOOKAHEAD}, output_scores=True) + {LOOKAHEAD}-position top-{TOPK} scan = the GPU arm's work; "
                      f"per-range mean seconds {[round(m, 2) for m in means]} weighted by the range sizes {[c for _, _, c in LENGTH_RANGES]}; "
                      f"{warmup} warm-up group(s); {cores} host threads")
    if with_full_call:
        p = prompt_of(111)
        t1 = time.time()
        scorer_oracle.get_yes_no_logprobs(model, tok, p, variant="100q", max_new_tokens=50)
        out["full_reference_call_50_s"] = round(time.time() - t1, 3)
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from llm_interpretation_replication_b200 import shapes
    r = cpu_reference_run(args.steps, args.warmup, with_full_call=True)
    cb = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"],
          "per_range_seconds": r["per_range_s"], "prefill_only_seconds_111_tokens": r["prefill_only_s"],
          "full_reference_call_50_tokens_seconds_111_tokens": r.get("full_reference_call_50_s")}
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * r["step_s"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(shapes.falcon_7b(), synthetic_lengths()),
            "cpu_baseline": cb,
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---- our arm ----------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--prompts", type=int, default=N_PROMPTS, help="prompts per step (default: the full 10,000)")
    ap.add_argument("--layers", type=int, default=0, help="debug only: truncate the model (result is then not a bench value)")
    ap.add_argument("--batch-tokens", type=int, default=262144,
                    help="packed prompt tokens per C-ABI call; 262144 = 2350 prompts per batch, which fills the waves of the one-token GEMMs (profiles/r1_bench_batch_tokens.jsonl)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--mode", default="fp16", choices=["bf16", "fp16", "bf16_split", "int8_emu"],
                    help="operand format of the tensor-core path: fp16 (weights converted from bf16; default), bf16 (fast, "
                         "misses the 2e-3 gate on some prompts) or bf16_split (bf16 weights, hi+lo bf16 activations)")
    ap.add_argument("--no-other-modes", action="store_true", help="skip the short BF16_FAST / BF16_SPLIT / 50-token legs (N = 1)")
    ap.add_argument("--resolve-rate", type=float, default=0.0,
                    help="not the headline workload: make this fraction of the prompts resolve at look-ahead position 0 "
                         "(candidate = the model's own greedy token), as real checkpoints mostly do; shows what compaction saves")
    ap.add_argument("--option", action="append", default=[], metavar="NAME=VALUE",
                    help="A/B only: b2s_set_option(NAME, VALUE) before the run (recorded in config.options)")
    ap.add_argument("--gemm-cta-group", type=int, default=-1,
                    help="tuning: 0 = per-launch policy (library default), 1 = one CTA per tile, 2 = CTA pairs, 3 = CTA pairs on 512x256 tiles")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch.distributed as dist
    from llm_interpretation_replication_b200 import _lib, shapes, weights
    from llm_interpretation_replication_b200.scorer import Scorer, plan_batches
    from llm_interpretation_replication_b200.dist import shard_prompts, gather_rows

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback for the b200 arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device(f"cuda:{local_rank}")
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    shp = shapes.falcon_7b()
    if args.layers:
        shp = shp.with_layers(args.layers)
    ids, cand, L = synthetic_prompts(shp, args.prompts)
    n_total = len(ids)
    mine = shard_prompts(L.tolist(), world)[rank]
    my_ids = [ids[i] for i in mine]
    my_cand = cand[mine]
    my_len = [int(L[i]) for i in mine]

    W = weights.init_weights_device(shp, seed=0, device=dev, weight_dtype=torch.bfloat16)
    max_seqs = max(2048, args.batch_tokens // 64)   # prompts per batch are bounded by tokens, not by this
    MODES = {"fp16": shapes.MODE_FP16, "bf16": shapes.MODE_BF16_FAST, "bf16_split": shapes.MODE_BF16_SPLIT,
             "int8_emu": shapes.MODE_INT8_EMU}
    mode = MODES[args.mode]
    sc = Scorer(shp, W, mode=mode, device=dev, max_tokens=args.batch_tokens, max_seqs=max_seqs,
                max_lookahead=LOOKAHEAD)
    lib = sc.lib
    if args.gemm_cta_group >= 0:
        _lib.check(lib.b2s_set_option(b"gemm_cta_group", args.gemm_cta_group), "set_option")
    for kv in args.option:
        name, val = kv.split("=")
        _lib.check(lib.b2s_set_option(name.encode(), int(val)), "set_option")
    batches = plan_batches(my_len, args.batch_tokens, max_seqs)
    if args.resolve_rate > 0:
        # position-0 pass to learn each prompt's greedy token, then make it candidate a for the chosen fraction
        _, g0 = sc.score_ids(my_ids, my_cand, topk=TOPK, max_lookahead=1)
        pick = np.random.RandomState(7).rand(len(my_ids)) < args.resolve_rate
        my_cand = my_cand.copy()
        my_cand[pick, 0] = g0[pick, 0]

    # host (pinned) and device-resident copies of every batch
    host_b, dev_b = [], []
    h2d = d2h = 0
    for b in batches:
        t_tok, t_cu, t_cand, mx = Scorer.pack([my_ids[i] for i in b], my_cand[b])
        o_h = torch.empty((len(b), 8), dtype=torch.float32).pin_memory()
        host_b.append((t_tok, t_cu, t_cand, o_h))
        dev_b.append((t_tok.to(dev), t_cu.to(dev), t_cand.to(dev), len(b), int(t_tok.numel()), mx,
                      torch.empty((len(b), 8), dtype=torch.float32, device=dev)))
        h2d += 4 * (t_tok.numel() + t_cu.numel() + t_cand.numel())
        d2h += 32 * len(b)
    local_rows = torch.empty((len(mine), 8), dtype=torch.float32, device=dev)

    def step_device():
        for (d_tok, d_cu, d_cand, n, tot, mx, d_out), b in zip(dev_b, batches):
            sc.score_batch_device(d_tok, d_cu, d_cand, n, tot, mx, TOPK, LOOKAHEAD, shp.eos_id, d_out, None)
            local_rows[torch.as_tensor(b, device=dev)] = d_out
        if world > 1:
            return gather_rows(local_rows, mine, n_total, world)
        return local_rows

    def step_host():
        for (t_tok, t_cu, t_cand, o_h), b in zip(host_b, batches):
            sc.score_batch_host(t_tok, t_cu, t_cand, TOPK, LOOKAHEAD, shp.eos_id, o_h, None)
            local_rows[torch.as_tensor(b, device=dev)] = o_h.to(dev, non_blocking=True)
        if world > 1:
            full = gather_rows(local_rows, mine, n_total, world)
        else:
            full = local_rows
        return full[:, 2].sum().item()     # device -> host read of the step's result

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms

    for _ in range(args.warmup):
        step_device()
    barrier()
    lib.b2s_launch_count(1)
    lib.b2s_profile_gemm(1)
    clocks = ClockSampler(local_rank) if rank == 0 else None
    ms = timed(step_device, args.steps)
    clk = clocks.stop() if clocks else {}
    lib.b2s_profile_gemm(0)
    g_ms, g_n, g_fl = C.c_double(), C.c_int64(), C.c_double()
    _lib.check(lib.b2s_profile_gemm_read(C.byref(g_ms), C.byref(g_n), C.byref(g_fl)), "profile read")
    launches = int(lib.b2s_launch_count(0))
    result = step_device()
    rel = result[:, 2]
    finite = bool(torch.isfinite(rel).all().item())
    value = n_total * args.steps / (ms / 1000.0)

    e2e = None
    if not args.no_e2e:
        step_host()
        ke = min(args.steps, 2)
        ms_e = timed(step_host, ke)
        e2e = {"value": n_total * ke / (ms_e / 1000.0), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h) + 4,
               "steps": ke, "api": "b2s_score_host (C-ABI, pinned host buffers)"}

    # ---- N = 1 extras: the same step in the other operand modes, and the entry points' 50-token call ----------------
    other, complete50 = {}, None
    if world == 1 and not args.no_other_modes and not args.layers and args.resolve_rate == 0:
        def one_mode(m):
            s2 = Scorer(shp, W, mode=MODES[m], device=dev, max_tokens=args.batch_tokens, max_seqs=max_seqs, max_lookahead=LOOKAHEAD)
            def stp():
                for (d_tok, d_cu, d_cand, n, tot, mx, d_out) in dev_b:
                    s2.score_batch_device(d_tok, d_cu, d_cand, n, tot, mx, TOPK, LOOKAHEAD, shp.eos_id, d_out, None)
            stp()
            ms2 = timed(stp, 1)
            fin = bool(torch.isfinite(dev_b[0][6][:, 2]).all().item())
            s2.close()
            del s2
            torch.cuda.empty_cache()
            return {"value": n_total / (ms2 / 1000.0), "unit": UNIT, "ms_per_step": ms2, "steps": 1, "warmup": 1, "all_rel_prob_finite": fin}
        for m in ("bf16", "bf16_split", "fp16"):
            if m != args.mode:
                other[m] = one_mode(m)
        # what the kept entry points call: score + complete to the reference's max_new_tokens = 50 (row N1)
        sc50 = Scorer(shp, W, mode=mode, device=dev, max_tokens=args.batch_tokens, max_seqs=max_seqs, max_lookahead=50)
        g_h = [torch.empty((len(b), 50), dtype=torch.int32).pin_memory() for b in batches]
        def step50():
            for (t_tok, t_cu, t_cand, o_h), gh in zip(host_b, g_h):
                sc50.score_batch_host(t_tok, t_cu, t_cand, TOPK, LOOKAHEAD, shp.eos_id, o_h, gh, gen_tokens=50)
        ms50 = timed(step50, 1)
        complete50 = {"value": n_total / (ms50 / 1000.0), "unit": UNIT, "ms_per_step": ms50, "steps": 1, "warmup": 0,
                      "api": "b2s_score_complete_host (scan 10 positions, generate 50 tokens per prompt, completions copied back)",
                      "d2h_bytes_per_step": int(d2h) + 200 * n_total}
        sc50.close()
        del sc50
        torch.cuda.empty_cache()

    if rank == 0:
        peak, peak_src = measured_peaks()
        achieved = g_fl.value / (g_ms.value / 1000.0) / 1e12 if g_ms.value > 0 else 0.0
        tokens = int(sum(my_len)) if world == 1 else int(L.sum())
        traffic = None
        tp = os.path.join(ROOT, "profiles", "gemm_traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("dram_bytes_per_launch")
        cfg = workload_config(shp, L)
        cfg.update({"batch_tokens": args.batch_tokens, "gemm_cta_group": max(args.gemm_cta_group, 0), "options": args.option,
                    "batches_per_rank": len(batches), "parallelism": f"prompt-sharded x{world}, replicated weights",
                    "l2": "inputs larger than L2: 13.8 GB of 16-bit weights are re-read for every batch",
                    "operand_mode": args.mode, "all_rel_prob_finite": finite})
        if args.resolve_rate > 0:
            cfg["resolve_rate"] = args.resolve_rate
            cfg["workload"] += f" -- VARIANT: {args.resolve_rate:.0%} of the prompts resolve at position 0 (not the headline workload)"
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": {"fp16": "fp16", "bf16": "bf16", "bf16_split": "bf16 (hi+lo activations)",
                          "int8_emu": "fp16 holding int8-grid values (LLM.int8 emulation)"}[args.mode], "data": "synthetic",
                "config": cfg,
                "clocks": clk, "gpu_launches": launches,
                "roofline": {"bound": "tensor", "kernel": "gemm_tc3_kernel / gemm_tc2_kernel (tcgen05 cta_group::2 on 512x256 / 256x256 pair tiles; gemm_tc_kernel when forced)", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                             "frac": achieved / peak if peak else None, "traffic": traffic, "peak_source": peak_src,
                             "traffic_source": "profiles/gemm_traffic.json (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, mean over the launches of one layer)",
                             "launches": int(g_n.value), "gemm_ms_per_step": g_ms.value / args.steps,
                             "gemm_share_of_step": (g_ms.value / ms) if ms else None,
                             "model_flops_frac": (shp.flops_per_token() * int(L.sum()) / world * args.steps / (ms / 1000.0) / 1e12) / peak},
                }
        if other:
            line["other_modes"] = other
        if complete50:
            line["complete50"] = complete50
        if e2e:
            line["e2e"] = e2e
        if world == 1 and not args.no_cpu_baseline:
            del sc, W
            torch.cuda.empty_cache()
            r = cpu_reference_run(1, 1)
            line["cpu_baseline"] = {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": r["sample"],
                                    "per_range_seconds": r["per_range_s"], "prefill_only_seconds_111_tokens": r["prefill_only_s"]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
