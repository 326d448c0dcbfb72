#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary8.txt
run() { name=$1; shift; timeout ${TMO:-600} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary8.txt; }
run t_gemm8      python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "gemm_tc"
run t_entry8     python -m pytest tests/test_gpu_entrypoints.py -q -m gpu
if grep -q "t_gemm8 exit 0" gpurun_out/summary8.txt; then
run gemm_bench8  python tools/gemm_bench.py 32768
run gemm_bench8b python tools/gemm_bench.py 111104
fi
tail -n 5 gpurun_out/t_gemm8.log; tail -n 25 gpurun_out/t_entry8.log; cat gpurun_out/gemm_bench8.log gpurun_out/gemm_bench8b.log
