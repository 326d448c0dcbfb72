#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary27.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary27.txt; }
run t_corr27 python -m pytest tests/test_gpu_corr.py tests/test_gpu_kernels.py -q -m gpu -k "corr or kappa or simt_kv"
run corrbench27 python tools/corr_bench.py
tail -n 30 gpurun_out/t_corr27.log; cat gpurun_out/corrbench27.log
