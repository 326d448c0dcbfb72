#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary4.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary4.txt; }
run t_quick4     python -m pytest tests/test_gpu_kernels.py tests/test_gpu_scorer.py -q -m gpu -k "attention or bf16 or edge"
run gemm_plain   python tools/gemm_bench.py 32768 ncu
if grep -q "gemm_plain exit 0" gpurun_out/summary4.txt; then
  TMO=900 run ncu_gemm ncu --set full --clock-control none --import-source on -k regex:gemm_tc -c 6 -o gpurun_out/gemm_full_r1 -f python tools/gemm_bench.py 32768 ncu
fi
run bench_1k4    python bench.py --prompts 1000 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline
if grep -q "bench_1k4 exit 0" gpurun_out/summary4.txt; then
  TMO=1200 run ncu_list4 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_r1b.csv python bench.py --prompts 1000 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline
fi
tail -n 5 gpurun_out/t_quick4.log gpurun_out/bench_1k4.log
