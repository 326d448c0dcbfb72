#!/bin/bash
# round 2, call G: tensor-core prompt-pass attention at head_dim 128 / with ALiBi, fused SwiGLU -- kernel tests, scorer goldens and
# real widths, the full-depth configs (C2 = StableLM-alpha-7B takes the new path)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_g.txt gpurun_out/full_depth_parity.jsonl
note() { echo "$*" | tee -a gpurun_out/summary_g.txt; }
timeout 600 python -m pytest tests/test_gpu_kernels.py -q --maxfail=10 -rf -k "attention" > gpurun_out/g_attn.log 2>&1; note "attention tests exit $?"
timeout 900 python -m pytest tests/test_gpu_scorer.py tests/test_gpu_checkpoint.py tests/test_gpu_entrypoints.py -q --maxfail=10 -rf > gpurun_out/g_scorer.log 2>&1; note "scorer tests exit $?"
timeout 900 python -m pytest tests/test_gpu_full_depth.py -q -rf > gpurun_out/g_depth.log 2>&1; note "full depth exit $?"
tail -n 5 gpurun_out/g_attn.log; tail -n 6 gpurun_out/g_scorer.log; tail -n 4 gpurun_out/g_depth.log; cat gpurun_out/full_depth_parity.jsonl 2>/dev/null | cut -c1-260
