#!/bin/bash
# round 2, call INT8: the LLM.int8 emulation mode -- fake-quantisation kernel, scorer vs the restated Linear8bitLt, a bench line in that mode
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_int8.txt gpurun_out/int8_parity.jsonl
note() { echo "$*" | tee -a gpurun_out/summary_int8.txt; }
timeout 900 python -m pytest tests/test_gpu_int8.py -q --maxfail=10 -rf > gpurun_out/int8_tests.log 2>&1; note "int8 tests exit $?"
tail -n 30 gpurun_out/int8_tests.log | cut -c1-300; cat gpurun_out/int8_parity.jsonl
timeout 900 python bench.py --mode int8_emu --no-other-modes --steps 1 --warmup 3 > gpurun_out/int8_bench.log 2>&1; note "bench int8_emu exit $?"
grep '^{' gpurun_out/int8_bench.log | tail -n 1 | cut -c1-700
