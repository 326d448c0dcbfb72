#!/bin/bash
# round 2, call K: launch list of one full-size Llama-2-7B pass (multi-head, head_dim 128: the most common shape of the sweep)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python tools/family_bench.py --models llama2_7b --prompts 600 > gpurun_out/k_plain.jsonl 2> gpurun_out/k_plain.err; echo "plain exit $?"
if grep -q prompts_per_s gpurun_out/k_plain.jsonl; then
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/llama_launches.csv \
    python tools/family_bench.py --models llama2_7b --prompts 600 > gpurun_out/k_ncu.log 2>&1; echo "ncu exit $?"
fi
cat gpurun_out/k_plain.jsonl | cut -c1-300; wc -l gpurun_out/llama_launches.csv
