#!/bin/bash
# round-end consistency pass without the ncu captures: full GPU suite, smoke, both bench arms with default flags
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_final2.txt
run() { name=$1; shift; timeout ${TMO:-1500} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_final2.txt; }
run g_tests   python -m pytest tests -x -q -m gpu
run g_smoke   python __graft_entry__.py smoke
run g_ref     python bench.py --impl reference --steps 3 --warmup 3
run g_bench   python bench.py
tail -n 3 gpurun_out/g_tests.log | cut -c1-200; tail -n 1 gpurun_out/g_smoke.log; tail -n 1 gpurun_out/g_bench.log | cut -c1-1500
