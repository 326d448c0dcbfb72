#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary9.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary9.txt; }
run t_all9       python -m pytest tests -x -q -m gpu
run bench9a      python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e
run bench9b      python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --batch-tokens 262144
tail -n 4 gpurun_out/t_all9.log; cat gpurun_out/bench9a.log gpurun_out/bench9b.log | cut -c1-1500
