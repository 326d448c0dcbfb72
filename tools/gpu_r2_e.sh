#!/bin/bash
# round 2, call E: resample kernel with more loads in flight -- tests, bench, and an ncu --set full capture of it at B = 10^7
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_e.txt
note() { echo "$*" | tee -a gpurun_out/summary_e.txt; }
python -m pytest tests/test_gpu_bootstrap.py tests/test_gpu_agreement.py -q --maxfail=5 -rf > gpurun_out/e_tests.log 2>&1; note "tests exit $?"
python tools/boot_bench.py > gpurun_out/e_boot.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"boot_resample_tile" -s 20 -c 2 -f -o gpurun_out/r2_resample_full python tools/boot_bench.py > gpurun_out/e_ncu.log 2>&1
note "ncu exit $?"
tail -n 3 gpurun_out/e_tests.log; tail -n 1 gpurun_out/e_boot.log
