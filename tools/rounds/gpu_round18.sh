#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary18.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary18.txt; }
run t_gemm18   python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "gemm_tc or lm_head"
run t_scorer18 python -m pytest tests/test_gpu_scorer.py -x -q -m gpu
if grep -q "t_gemm18 exit 0" gpurun_out/summary18.txt; then
run sweep18    python tools/sweep_gemm.py 111104 1.5
run bench18    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
fi
tail -n 4 gpurun_out/t_gemm18.log; tail -n 4 gpurun_out/t_scorer18.log; cat gpurun_out/sweep18.log; tail -n 1 gpurun_out/bench18.log | cut -c1-200; tail -n 1 gpurun_out/bench18.log | grep -o '"roofline.*' | cut -c1-300
