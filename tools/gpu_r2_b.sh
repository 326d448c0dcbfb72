#!/bin/bash
# round 2, call B: whole GPU suite again (fixed tests, BF16_SPLIT with fp16 attention, new bootstrap kernels, entry points 6/7),
# bootstrap bench, 2000-prompt parity histogram
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_b.txt gpurun_out/full_depth_parity.jsonl gpurun_out/golden_position_mismatches.jsonl
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_b.txt; }
TMO=1200 run b_tests python -m pytest tests -q -m gpu --maxfail=12 -rf --durations=8
run b_boot python tools/boot_bench.py
TMO=700 run b_hist python tools/parity_histogram.py ${HIST_N:-2000}
tail -n 25 gpurun_out/b_tests.log; tail -n 2 gpurun_out/b_boot.log; grep -v '^ \|^{$\|^}$' gpurun_out/b_hist.log | cut -c1-700
cat gpurun_out/full_depth_parity.jsonl 2>/dev/null
