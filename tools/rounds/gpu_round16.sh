#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary16.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary16.txt; }
run t_gemm16   python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "gemm_tc or lm_head"
run t_scorer16 python -m pytest tests/test_gpu_scorer.py -x -q -m gpu
if grep -q "t_gemm16 exit 0" gpurun_out/summary16.txt; then
run sweep16    python tools/sweep_gemm.py 111104 1.5
fi
tail -n 12 gpurun_out/t_gemm16.log; tail -n 6 gpurun_out/t_scorer16.log; cat gpurun_out/sweep16.log
