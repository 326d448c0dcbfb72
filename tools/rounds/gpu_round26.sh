#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary26.txt
run() { name=$1; shift; timeout ${TMO:-1200} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary26.txt; }
run t_all26 python -m pytest tests -q -m gpu
run bench26 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
tail -n 25 gpurun_out/t_all26.log; tail -n 1 gpurun_out/bench26.log | cut -c1-200
