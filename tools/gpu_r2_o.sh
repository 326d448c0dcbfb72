#!/bin/bash
# round 2, call O: the BF16_SPLIT_ATTN32 mode: scorer tests in every mode, then Llama-2-7B at full depth against the fp32 check mode
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests/test_gpu_scorer.py -q --maxfail=10 -rf > gpurun_out/o_scorer.log 2>&1; echo "scorer tests exit $?"
tail -n 6 gpurun_out/o_scorer.log | cut -c1-250
timeout 900 python tools/parity_histogram.py 400 0 bf16_split_attn32,fp16,bf16_split llama2_7b > gpurun_out/parity_llama.json 2> gpurun_out/parity_llama.err; echo "histogram exit $?"
tail -n 4 gpurun_out/parity_llama.err | cut -c1-420
