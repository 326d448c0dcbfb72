#!/bin/bash
# round 2, call F: SwiGLU fusion tests (Llama family), then compute-sanitizer memcheck over a subset of the GPU suite
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_f.txt
note() { echo "$*" | tee -a gpurun_out/summary_f.txt; }
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_scorer.py tests/test_gpu_checkpoint.py -q --maxfail=10 -rf > gpurun_out/f_tests.log 2>&1; note "tests exit $?"
timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 3 python -m pytest tests/test_gpu_scorer.py tests/test_gpu_bootstrap.py tests/test_gpu_agreement.py -q -x \
  -k "(golden and (0 or 3 or 5)) or edge_cases or range_overflow or stream_continues or kernels_match or philox_indices or membership" > gpurun_out/f_memcheck.log 2>&1
note "memcheck exit $?"
tail -n 4 gpurun_out/f_tests.log; grep -E "ERROR SUMMARY|passed|failed" gpurun_out/f_memcheck.log | tail -4
