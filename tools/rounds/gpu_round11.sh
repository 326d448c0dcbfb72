#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary11.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary11.txt; }
run parity_full  python tools/full_size_parity.py 16 32
run bench_1k11   python bench.py --prompts 1000 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline
if grep -q "bench_1k11 exit 0" gpurun_out/summary11.txt; then
  TMO=1200 run ncu_list11 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_r1c.csv python bench.py --prompts 1000 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline
fi
tail -n 20 gpurun_out/parity_full.log
