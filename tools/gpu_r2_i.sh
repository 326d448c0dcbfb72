#!/bin/bash
# round 2, call I: streaming one-token attention for head_dim 80 / 128 / ALiBi / few-heads-per-KV-head models: A/B, then the GPU suite
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_i.txt
note() { echo "$*" | tee -a gpurun_out/summary_i.txt; }
timeout 900 python tools/step_attn_ab.py > gpurun_out/step_attn_ab.jsonl 2> gpurun_out/step_attn_ab.err; note "A/B exit $?"
timeout 2400 python -m pytest tests -q -m gpu --maxfail=15 -rf > gpurun_out/i_tests.log 2>&1; note "gpu suite exit $?"
cat gpurun_out/step_attn_ab.jsonl; tail -n 3 gpurun_out/step_attn_ab.err; tail -n 12 gpurun_out/i_tests.log | cut -c1-250
