#!/bin/bash
# round 2, call J: head_dim 80 on the tensor-core prompt-pass kernel (3D TMA zero fill): kernel test, RedPajama-3B real widths,
# C1 at full depth, then the remaining GPU suite
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_j.txt
note() { echo "$*" | tee -a gpurun_out/summary_j.txt; }
timeout 600 python -m pytest tests/test_gpu_kernels.py -q -k "attention" --maxfail=10 -rf > gpurun_out/j_kern.log 2>&1; note "attention kernel tests exit $?"
timeout 900 python -m pytest tests/test_gpu_scorer.py -q -k "redpajama or neox or overflow or golden" --maxfail=10 -rf > gpurun_out/j_scorer.log 2>&1; note "scorer tests exit $?"
timeout 900 python -m pytest tests/test_gpu_full_depth.py -q -k "c1" -rf > gpurun_out/j_c1.log 2>&1; note "C1 full depth exit $?"
tail -n 8 gpurun_out/j_kern.log | cut -c1-300; tail -n 6 gpurun_out/j_scorer.log | cut -c1-300; tail -n 4 gpurun_out/j_c1.log | cut -c1-300
