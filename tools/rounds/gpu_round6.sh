#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary6.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary6.txt; }
run t_attn6      python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "attention"
run t_scorer6    python -m pytest tests/test_gpu_scorer.py -q -m gpu
run bench_quick6 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
tail -n 15 gpurun_out/t_attn6.log; tail -n 5 gpurun_out/t_scorer6.log gpurun_out/bench_quick6.log
