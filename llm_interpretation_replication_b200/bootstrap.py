"""Host-side mirror of the reference's bootstrap functions on top of libb200score.

Kept names, arguments and return dicts:
  * bootstrap_mae_difference, bootstrap_overall_mae_difference
      -- reference survey_analysis/analyze_base_vs_instruct_mae_100q.py:289-345, 348-414
  * analyze_results_bootstrap (the resampling core of analyze_results)
      -- reference analysis/run_base_vs_instruct_100q.py:614-705
  * prompt_bootstrap_and_signflip
      -- reference analysis/analyze_results_base_versus_instruct.py:300-365
The `for _ in range(n_bootstrap)` loops become: one MT19937 kernel that reproduces NumPy's legacy index
stream bit for bit, one gather/mean kernel, one sort + percentile kernel.  Only scalars come back.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import numpy as np
import torch

from . import _lib


class NumpyLegacyStream:
    """Device-resident equivalent of ``np.random.seed(seed)`` + repeated ``np.random.choice``."""

    def __init__(self, seed: int, device=None):
        if not torch.cuda.is_available():
            raise _lib.B2SError("no CUDA device: the bootstrap kernels have no CPU fallback")
        self.lib = _lib.load()
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.state = torch.empty(625, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2s_mt19937_seed(int(seed) & 0xFFFFFFFF, _lib.ptr(self.state), _lib.current_stream()),
                       "b2s_mt19937_seed")

    def choice(self, n: int, count: int) -> torch.Tensor:
        """``count`` draws of ``np.random.choice(n, ...)`` in stream order (int32, device)."""
        out = torch.empty(int(count), dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2s_mt19937_choice(_lib.ptr(self.state), int(n), int(count), _lib.ptr(out),
                                                   _lib.current_stream()), "b2s_mt19937_choice")
        return out

    def choice_into(self, n: int, out: torch.Tensor) -> None:
        """``np.random.choice(n, out.numel())`` written into a contiguous int32 device view (no allocation)."""
        assert out.dtype == torch.int32 and out.is_contiguous() and out.device == self.device
        with torch.cuda.device(self.device):
            _lib.check(self.lib.b2s_mt19937_choice(_lib.ptr(self.state), int(n), int(out.numel()), _lib.ptr(out),
                                                   _lib.current_stream()), "b2s_mt19937_choice")


def resample(vals: np.ndarray, split: int, B: int, idx: Optional[torch.Tensor] = None, philox_seed: int = 0,
             device=None) -> torch.Tensor:
    """boot[B,3] = (mean rows<split, mean rows>=split, mean|.| all rows) for B resamples of the columns."""
    lib = _lib.load()
    device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
    v = torch.as_tensor(np.ascontiguousarray(vals, dtype=np.float64)).to(device)
    F, n = v.shape
    boot = torch.empty((B, 3), dtype=torch.float64, device=device)
    if idx is not None:
        assert idx.dtype == torch.int32 and idx.numel() == B * n and idx.is_cuda
    with torch.cuda.device(device):
        _lib.check(lib.b2s_bootstrap_resample(_lib.ptr(v), F, split, n, B, _lib.ptr(idx), int(philox_seed),
                                              _lib.ptr(boot), _lib.current_stream()), "b2s_bootstrap_resample")
    return boot


def summary(boot: torch.Tensor, stat: int, q_lo: float, q_hi: float, observed: float, want_stat: bool = False):
    """(dict, stat vector or None).  dict keys: pct_lo, pct_hi, frac_le0, frac_ge0, mean, frac_ge_obs, min, max."""
    lib = _lib.load()
    B = boot.shape[0]
    dev = boot.device
    nbytes = lib.b2s_bootstrap_scratch_bytes(B)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    out = torch.empty(8, dtype=torch.float64, device=dev)
    stat_out = torch.empty(B, dtype=torch.float64, device=dev) if want_stat else None
    with torch.cuda.device(dev):
        _lib.check(lib.b2s_bootstrap_summary(_lib.ptr(boot), B, stat, float(q_lo), float(q_hi), float(observed),
                                             _lib.ptr(out), _lib.ptr(stat_out), _lib.ptr(scratch), nbytes,
                                             _lib.current_stream()), "b2s_bootstrap_summary")
    o = out.cpu().numpy()
    keys = ("pct_lo", "pct_hi", "frac_le0", "frac_ge0", "mean", "frac_ge_obs", "min", "max")
    return dict(zip(keys, map(float, o))), stat_out


def _two_sided_p(observed: float, s: dict, cap: bool = True) -> float:
    p = 2 * s["frac_le0"] if observed > 0 else 2 * s["frac_ge0"]
    return min(p, 1.0) if cap else p


N_BOOTSTRAP = 10000
RANDOM_SEED = 42
CONFIDENCE_LEVEL = 0.95


def bootstrap_mae_difference(base_errors, instruct_errors, n_bootstrap=N_BOOTSTRAP, seed=RANDOM_SEED,
                             stream: Optional[NumpyLegacyStream] = None, device=None):
    """Same contract as the reference function (dict incl. ``bootstrap_diffs``)."""
    base_mae = np.mean(base_errors)
    instruct_mae = np.mean(instruct_errors)
    observed_diff = instruct_mae - base_mae
    n = min(len(base_errors), len(instruct_errors))
    vals = np.stack([np.asarray(base_errors[:n], dtype=np.float64), np.asarray(instruct_errors[:n], dtype=np.float64)])
    rs = stream if stream is not None else NumpyLegacyStream(seed, device)
    idx = rs.choice(n, n_bootstrap * n)
    boot = resample(vals, 1, n_bootstrap, idx, device=rs.device)
    alpha = 1 - CONFIDENCE_LEVEL
    s, diffs = summary(boot, 1, (alpha / 2) * 100, (1 - alpha / 2) * 100, observed_diff, want_stat=True)
    return dict(base_mae=base_mae, instruct_mae=instruct_mae, observed_diff=observed_diff, ci_lower=s["pct_lo"],
                ci_upper=s["pct_hi"], p_value=_two_sided_p(observed_diff, s), n_questions=n,
                bootstrap_diffs=diffs.cpu().numpy())


def bootstrap_overall_mae_difference(all_family_errors: Dict[str, dict], n_bootstrap=N_BOOTSTRAP, seed=RANDOM_SEED,
                                     device=None):
    n = None
    for fam, e in all_family_errors.items():
        m = min(len(e["base"]), len(e["instruct"]))
        n = m if n is None else min(n, m)
    fams = list(all_family_errors.keys())
    Bm = np.array([all_family_errors[f]["base"][:n] for f in fams], dtype=np.float64)
    Im = np.array([all_family_errors[f]["instruct"][:n] for f in fams], dtype=np.float64)
    obs_b, obs_i = np.mean(Bm), np.mean(Im)
    observed_diff = obs_i - obs_b
    rs = NumpyLegacyStream(seed, device)
    idx = rs.choice(n, n_bootstrap * n)
    boot = resample(np.concatenate([Bm, Im], axis=0), len(fams), n_bootstrap, idx, device=rs.device)
    alpha = 1 - CONFIDENCE_LEVEL
    s, diffs = summary(boot, 1, (alpha / 2) * 100, (1 - alpha / 2) * 100, observed_diff, want_stat=True)
    return dict(base_mae=obs_b, instruct_mae=obs_i, observed_diff=observed_diff, ci_lower=s["pct_lo"],
                ci_upper=s["pct_hi"], p_value=_two_sided_p(observed_diff, s), n_questions=n, n_models=len(fams),
                bootstrap_diffs=diffs.cpu().numpy())


def analyze_results_bootstrap(family_diffs: Dict[str, np.ndarray], n_bootstrap=N_BOOTSTRAP, seed=RANDOM_SEED, device=None):
    """Resampling core of the reference's analyze_results(): ONE seed, then the families in order, then the
    concatenation, all on the same stream (run_base_vs_instruct_100q.py:616,655-658,692-695)."""
    rs = NumpyLegacyStream(seed, device)
    out, kept = {}, {}
    for fam, diffs in family_diffs.items():
        diffs = np.asarray(diffs, dtype=np.float64)
        if len(diffs) < 10:
            continue
        kept[fam] = diffs
        n = len(diffs)
        obs = float(np.mean(diffs))
        boot = resample(diffs[None, :], 1, n_bootstrap, rs.choice(n, n_bootstrap * n), device=rs.device)
        s, _ = summary(boot, 0, 2.5, 97.5, obs)
        out[fam] = dict(n_prompts=n, mean_diff=obs, mae=float(np.mean(np.abs(diffs))), ci_lower=s["pct_lo"],
                        ci_upper=s["pct_hi"], p_value=_two_sided_p(obs, s))
    overall = None
    if len(kept) >= 2:
        alld = np.concatenate(list(kept.values()))
        n = len(alld)
        obs = float(np.mean(alld))
        boot = resample(alld[None, :], 1, n_bootstrap, rs.choice(n, n_bootstrap * n), device=rs.device)
        s, _ = summary(boot, 0, 2.5, 97.5, obs)
        overall = dict(mean_diff=obs, mae=float(np.mean(np.abs(alld))), ci_lower=s["pct_lo"], ci_upper=s["pct_hi"],
                       p_value=_two_sided_p(obs, s))
    return out, overall


def prompt_bootstrap_and_signflip(pivot: np.ndarray, n_bootstrap=N_BOOTSTRAP, seed=RANDOM_SEED, device=None,
                                  observed_mae: Optional[float] = None, observed_mean: Optional[float] = None):
    """reference analyze_results_base_versus_instruct.py:300-365.  pivot [n_prompts, F]; rows are resampled,
    statistics are taken over the flattened resample.  ``observed_mae`` / ``observed_mean`` are what the script compares
    against: pandas means over the UN-pivoted frame (``prompt_diff_df['Abs_Difference'].mean()``, :327-328), whose rows
    and summation order differ from the pivot's; default = the pivot's own means.
    The sign-flip "permutation" MAE is ``np.mean(np.abs(x * s))`` = ``np.mean(np.abs(x))`` bit for bit for every draw
    (|x * s| == |x| exactly), so ``p_value_mae = mean(perm_maes >= observed_mae)`` (:350-357) is 1.0 or 0.0 according
    to that one comparison; the draws are still taken from the stream (n_bootstrap * n*F words) to keep its position
    identical to the reference's."""
    pivot = np.asarray(pivot, dtype=np.float64)
    n, F = pivot.shape
    perm_mae = float(np.mean(np.abs(pivot.flatten())))            # every entry of the script's perm_maes
    observed_mae = perm_mae if observed_mae is None else float(observed_mae)
    observed_mean = float(np.mean(pivot.flatten())) if observed_mean is None else float(observed_mean)
    rs = NumpyLegacyStream(seed, device)
    idx = rs.choice(n, n_bootstrap * n)
    boot = resample(np.ascontiguousarray(pivot.T), F, n_bootstrap, idx, device=rs.device)
    s_mae, _ = summary(boot, 2, 2.5, 97.5, observed_mae)
    s_mean, _ = summary(boot, 0, 2.5, 97.5, observed_mean)
    signs = rs.choice(2, n_bootstrap * n * F)          # choice([-1, 1]) consumes one masked word per draw
    del signs
    p_mean = 2 * s_mean["frac_le0"] if observed_mean > 0 else 2 * s_mean["frac_ge0"]
    return dict(mae=observed_mae, mae_ci_lower=s_mae["pct_lo"], mae_ci_upper=s_mae["pct_hi"],
                mae_p_value=1.0 if perm_mae >= observed_mae else 0.0,
                mean_diff=observed_mean, mean_ci_lower=s_mean["pct_lo"], mean_ci_upper=s_mean["pct_hi"],
                mean_p_value=p_mean, n_prompts=n, n_models=F)


# ---- simulated-individual correlation bootstrap (bootstrap_confidence_intervals.py) ----------------------------
CI_GROUPS = {g: [f"Q{g}_{i}" for i in range(1, 12) if i != 8] for g in range(1, 6)}   # reference :45-51


def build_ci_tables(model_responses: Dict[str, dict], question_mapping: Dict[str, str], by_question: Dict[str, dict]):
    """Flatten the script's dict lookups (:67-90) into dense tables for the kernel: for group g and its k-th
    question, the FIRST prompt of question_mapping that maps to the id (the script's linear search), whether that
    prompt is in the model's responses and the id has human statistics, the model's rel_prob, mean/std on 0-1."""
    models = list(model_responses.keys())
    M = len(models)
    avail = np.zeros((M, 5, 10), np.uint8)
    rel = np.full((M, 5, 10), np.nan, np.float64)
    mean = np.zeros((5, 10), np.float64)
    std = np.zeros((5, 10), np.float64)
    first_prompt = {}
    for p, q in question_mapping.items():
        first_prompt.setdefault(q, p)
    for g in range(1, 6):
        for k, qid in enumerate(CI_GROUPS[g]):
            p = first_prompt.get(qid)
            if qid in by_question:
                mean[g - 1, k] = by_question[qid]["mean_response"] / 100.0
                std[g - 1, k] = by_question[qid]["std_response"] / 100.0
            for m, name in enumerate(models):
                r = model_responses[name]
                if p and p in r and qid in by_question:
                    avail[m, g - 1, k] = 1
                    rel[m, g - 1, k] = r[p]["rel_prob"]
    types = np.array([0 if model_responses[n][next(iter(model_responses[n]))]["type"] == "base" else 1 for n in models],
                     dtype=np.int32)
    return models, avail, rel, mean, std, types


def simulate_individual_correlations(tables, n_samples: int, seed0: int, n_seeds: int, want_corrs: bool = False,
                                     want_state: bool = False, device=None):
    """means [n_seeds, 2] (base, instruct); optionally the first seed's per-sample correlations [M, n_samples]
    and the MT19937 state after the last seed (a NumpyLegacyStream that continues it)."""
    lib = _lib.load()
    if not torch.cuda.is_available():
        raise _lib.B2SError("no CUDA device: the bootstrap kernels have no CPU fallback")
    device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
    models, avail, rel, mean, std, types = tables
    d = lambda a: torch.as_tensor(np.ascontiguousarray(a)).to(device)
    t_av, t_rel, t_mean, t_std, t_ty = d(avail), d(rel), d(mean), d(std), d(types)
    means = torch.empty((n_seeds, 2), dtype=torch.float64, device=device)
    corrs = torch.empty((len(models), n_samples), dtype=torch.float64, device=device) if want_corrs else None
    stream_obj = None
    state = None
    if want_state:
        stream_obj = NumpyLegacyStream(0, device)
        state = stream_obj.state
    with torch.cuda.device(device):
        _lib.check(lib.b2s_ci_simulate(_lib.ptr(t_av), _lib.ptr(t_rel), _lib.ptr(t_mean), _lib.ptr(t_std), _lib.ptr(t_ty),
                                       len(models), n_samples, int(seed0) & 0xFFFFFFFF, n_seeds, _lib.ptr(means),
                                       _lib.ptr(corrs), _lib.ptr(state), _lib.current_stream()), "b2s_ci_simulate")
    return means, corrs, stream_obj


def percentile_ci(x: torch.Tensor, q_lo=2.5, q_hi=97.5):
    """np.percentile CI of a device vector (NaNs must have been removed) through the sort/percentile kernel."""
    boot = torch.zeros((x.numel(), 3), dtype=torch.float64, device=x.device)
    boot[:, 0] = x
    s, _ = summary(boot, 0, q_lo, q_hi, 0.0)
    return s["pct_lo"], s["pct_hi"]
