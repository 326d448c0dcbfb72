#!/bin/bash
# round 2, call L: vectorised RoPE + cache append, deeper unroll of the streaming one-token attention: parity tests, then throughput
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_l.txt
note() { echo "$*" | tee -a gpurun_out/summary_l.txt; }
timeout 1200 python -m pytest tests/test_gpu_scorer.py tests/test_gpu_checkpoint.py -q --maxfail=10 -rf > gpurun_out/l_scorer.log 2>&1; note "scorer tests exit $?"
timeout 900 python -m pytest tests/test_gpu_full_depth.py -q -rf > gpurun_out/l_depth.log 2>&1; note "full depth exit $?"
timeout 900 python tools/family_bench.py --models llama2_7b stablelm_alpha_7b mistral_7b redpajama_3b bloomz_7b1 > gpurun_out/l_family.jsonl 2> gpurun_out/l_family.err; note "family bench exit $?"
tail -n 5 gpurun_out/l_scorer.log | cut -c1-250; tail -n 4 gpurun_out/l_depth.log | cut -c1-250; cut -c1-250 gpurun_out/l_family.jsonl
