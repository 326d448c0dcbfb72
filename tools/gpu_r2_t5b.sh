#!/bin/bash
# round 2, call T5-b: encoder-decoder tests again (tied head, per-mode bounds), the entry points that now run T5 models,
# throughput at T0_3B / tk-instruct size and at t5-v1_1-base size
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_t5b.txt gpurun_out/t5_golden_parity.jsonl
note() { echo "$*" | tee -a gpurun_out/summary_t5b.txt; }
timeout 900 python -m pytest tests/test_gpu_t5.py tests/test_gpu_entrypoints.py tests/test_gpu_checkpoint.py -q --maxfail=20 -rf > gpurun_out/t5b_tests.log 2>&1; note "t5 + entry point tests exit $?"
timeout 900 python tools/t5_bench.py --prompts 2000 --max-tokens 65536 > gpurun_out/t5_bench_xl.jsonl 2> gpurun_out/t5_bench_xl.err; note "t5 bench xl exit $?"
timeout 600 python tools/t5_bench.py --shape t5-v1_1-base --prompts 4000 > gpurun_out/t5_bench_base.jsonl 2> gpurun_out/t5_bench_base.err; note "t5 bench base exit $?"
tail -n 15 gpurun_out/t5b_tests.log | cut -c1-300; cat gpurun_out/t5_bench_xl.jsonl gpurun_out/t5_bench_base.jsonl; tail -n 5 gpurun_out/t5_bench_xl.err
