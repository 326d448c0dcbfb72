#!/bin/bash
# round 2, call T5: encoder-decoder scorer -- goldens of the reference's is_encoder_decoder branch in all four modes, completions,
# compaction / batching bit-identity, t5-v1_1-base widths; then the kernels touched on the way (attention, SwiGLU epilogue)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_t5.txt gpurun_out/t5_golden_parity.jsonl
note() { echo "$*" | tee -a gpurun_out/summary_t5.txt; }
timeout 900 python -m pytest tests/test_gpu_t5.py -q --maxfail=20 -rf > gpurun_out/t5_tests.log 2>&1; note "t5 tests exit $?"
timeout 600 python -m pytest tests/test_gpu_kernels.py -q --maxfail=10 -rf -k "attention or swiglu or SwiGLU" > gpurun_out/t5_kern.log 2>&1; note "kernel tests exit $?"
timeout 600 python -m pytest tests/test_gpu_scorer.py -q --maxfail=10 -rf -k "golden" > gpurun_out/t5_scorer.log 2>&1; note "scorer goldens exit $?"
tail -n 40 gpurun_out/t5_tests.log | cut -c1-300; tail -n 4 gpurun_out/t5_kern.log; tail -n 4 gpurun_out/t5_scorer.log; cat gpurun_out/t5_golden_parity.jsonl 2>/dev/null
