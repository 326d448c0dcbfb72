#!/bin/bash
# round 2, call A: whole GPU suite with the new full-depth / range / compaction tests, a short bench, the parity histogram
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_a.txt gpurun_out/full_depth_parity.jsonl gpurun_out/golden_position_mismatches.jsonl
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_a.txt; }
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/a_gpu.txt; free -g >> gpurun_out/a_gpu.txt; nproc >> gpurun_out/a_gpu.txt
TMO=1200 run a_tests python -m pytest tests -q -m gpu --maxfail=12 -rf --durations=15
run a_bench python bench.py --steps 1 --warmup 1 --no-cpu-baseline
TMO=700 run a_hist python tools/parity_histogram.py ${HIST_N:-1000}
cp gpurun_out/a_hist.log gpurun_out/parity_histogram_raw.txt 2>/dev/null
tail -n 40 gpurun_out/a_tests.log; tail -n 1 gpurun_out/a_bench.log | cut -c1-3000; tail -n 5 gpurun_out/a_hist.log | cut -c1-1500
cat gpurun_out/full_depth_parity.jsonl 2>/dev/null
