#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary22.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary22.txt; }
TMO=300 run t_attn22  python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "attention"
if grep -q "t_attn22 exit 0" gpurun_out/summary22.txt; then
run attn22a python tools/attn_bench.py longest
run attn22b python tools/attn_bench.py typical
run t_scorer22 python -m pytest tests/test_gpu_scorer.py -x -q -m gpu
run bench22    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
fi
tail -n 15 gpurun_out/t_attn22.log; cat gpurun_out/attn22a.log gpurun_out/attn22b.log; tail -n 3 gpurun_out/t_scorer22.log; tail -n 1 gpurun_out/bench22.log | cut -c1-200
