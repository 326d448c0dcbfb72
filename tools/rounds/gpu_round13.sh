#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary13.txt
run() { name=$1; shift; timeout ${TMO:-1200} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary13.txt; }
run t_all13      python -m pytest tests -x -q -m gpu
run boot_bench13 python tools/boot_bench.py
run bench13      python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
tail -n 5 gpurun_out/t_all13.log; cat gpurun_out/boot_bench13.log; tail -n 1 gpurun_out/bench13.log | cut -c1-1600
