#!/bin/bash
# round 2, call T5-e: encoder-decoder tests and throughput after a kernel change (default batch size)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_t5e.txt gpurun_out/t5_golden_parity.jsonl gpurun_out/t5e_bench.jsonl
note() { echo "$*" | tee -a gpurun_out/summary_t5e.txt; }
timeout 900 python -m pytest tests/test_gpu_t5.py -q --maxfail=20 -rf > gpurun_out/t5e_tests.log 2>&1; note "t5 tests exit $?"
timeout 900 python tools/t5_bench.py --prompts 4800 --modes fp16_ffn_split fp16 bf16_split --steps 2 --max-tokens 131072 --max-seqs 2048 >> gpurun_out/t5e_bench.jsonl 2> gpurun_out/t5e_bench.err; note "bench xl exit $?"
timeout 600 python tools/t5_bench.py --shape t5-v1_1-base --prompts 6000 --modes fp16_ffn_split fp16 bf16_split --steps 2 --max-tokens 131072 --max-seqs 2048 >> gpurun_out/t5e_bench.jsonl 2>> gpurun_out/t5e_bench.err; note "bench base exit $?"
tail -n 6 gpurun_out/t5e_tests.log | cut -c1-300; cat gpurun_out/t5e_bench.jsonl; tail -n 3 gpurun_out/t5e_bench.err
