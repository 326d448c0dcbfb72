#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary15.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary15.txt; }
run t_gemm15   python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "gemm_tc or lm_head"
if grep -q "t_gemm15 exit 0" gpurun_out/summary15.txt; then
run sweep15    python tools/sweep_gemm.py 111104 1.5
fi
tail -n 6 gpurun_out/t_gemm15.log; cat gpurun_out/sweep15.log
