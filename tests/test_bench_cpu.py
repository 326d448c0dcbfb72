"""bench.py host logic (no GPU): the synthetic workload has the perturbation length histogram, both arms describe the same
workload, and the reference arm's line carries every key the contract names (run on a tiny stand-in shape)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_synthetic_workload_matches_the_perturbation_histogram():
    L = bench.synthetic_lengths()
    assert len(L) == 10000 and L.min() >= 66 and L.max() <= 190
    h = np.histogram(L, bins=[64, 96, 128, 160, 192])[0]
    assert h.tolist() == [2664, 5257, 1726, 353]
    assert int(L.sum()) == 1109479                             # the token count DESIGN.md / BASELINE.md quote
    assert [c for _, _, c in bench.LENGTH_RANGES] == h.tolist()


def test_both_arms_describe_the_same_workload(monkeypatch, capsys):
    from llm_interpretation_replication_b200 import shapes
    real = shapes.falcon_7b()
    cfg = bench.workload_config(real, bench.synthetic_lengths())
    assert cfg["workload"] == bench.WORKLOAD and cfg["n_prompts"] == 10000 and cfg["max_lookahead"] == 10 and cfg["topk"] == 5
    monkeypatch.setattr(shapes, "falcon_7b", lambda: shapes.tiny_falcon(vocab=512))   # the CPU arm on a stand-in that runs in seconds

    class A:
        steps, warmup, gpus = 1, 0, 1
    bench.run_reference_arm(A)
    line = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "cpu_baseline", "e2e", "gpu_launches"):
        assert k in line, k
    assert line["impl"] == "reference" and line["metric"] == bench.METRIC and line["unit"] == bench.UNIT and line["higher_is_better"]
    assert line["e2e"] == {"value": line["value"], "unit": bench.UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = line["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == line["value"] and "max_new_tokens=10" in cb["sample"]
    assert len(cb["per_range_seconds"]) == 4 and cb["full_reference_call_50_tokens_seconds_111_tokens"] > 0
    same = {k: v for k, v in line["config"].items() if k != "n_layer"}
    assert same == {k: v for k, v in cfg.items() if k != "n_layer"}       # (the stand-in only differs in depth)
