#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary7.txt
run() { name=$1; shift; timeout ${TMO:-600} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary7.txt; }
run t_gemm7      python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "gemm_tc or lm_head"
if grep -q "t_gemm7 exit 0" gpurun_out/summary7.txt; then
run gemm_bench7  python tools/gemm_bench.py 32768
run gemm_bench7b python tools/gemm_bench.py 111104
fi
tail -n 25 gpurun_out/t_gemm7.log; cat gpurun_out/gemm_bench7.log gpurun_out/gemm_bench7b.log
