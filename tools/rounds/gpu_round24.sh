#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary24.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary24.txt; }
TMO=300 run t_attn24  python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "attention"
if grep -q "t_attn24 exit 0" gpurun_out/summary24.txt; then
run dec24   python tools/dec_attn_bench.py
run t_all24 python -m pytest tests -x -q -m gpu
run bench24 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
fi
tail -n 15 gpurun_out/t_attn24.log; cat gpurun_out/dec24.log; tail -n 3 gpurun_out/t_all24.log; tail -n 1 gpurun_out/bench24.log | cut -c1-200
