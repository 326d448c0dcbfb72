#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary3.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary3.txt; }
run t_attn       python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "attention"
run t_scorerbf3  python -m pytest tests/test_gpu_scorer.py -q -m gpu -k "bf16"
run bench_quick3 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
tail -n 12 gpurun_out/t_attn.log; tail -n 5 gpurun_out/t_scorerbf3.log gpurun_out/bench_quick3.log
