#!/bin/bash
# round 2, call C: ncu evidence for the kernels the bench runs by default -- launch list of one bench step and --set full
# captures of the four prompt-pass GEMMs of a layer, LayerNorm, prompt-pass attention and the bootstrap kernels.
# Every ncu command follows a plain run of the same command line that exited 0.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_c.txt
B="python bench.py --prompts 1200 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --no-other-modes"
BOOT="python tools/boot_bench.py --small"
note() { echo "$*" | tee -a gpurun_out/summary_c.txt; }
python -m pytest tests/test_gpu_scorer.py tests/test_gpu_kernels.py -q -x -k "opt_bloom or golden or epilogues or default_policy" > gpurun_out/c_tests.log 2>&1; note "tests exit $?"
$B > gpurun_out/c_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/r2_launches.csv $B > gpurun_out/c_list.log 2>&1
note "launch list exit $?"
$B > gpurun_out/c_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"gemm_tc" -s 8 -c 4 -f -o gpurun_out/r2_gemm_full $B > gpurun_out/c_full_gemm.log 2>&1
note "gemm full exit $?"
$B > gpurun_out/c_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"ln_kernel|attn_tc2_kernel|lmhead_merge|scan_update" -s 4 -c 4 -f -o gpurun_out/r2_misc_full $B > gpurun_out/c_full_misc.log 2>&1
note "misc full exit $?"
$BOOT > gpurun_out/c_boot_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"boot_resample_tile|mt_gen|mt_scatter|mt_count" -c 6 -f -o gpurun_out/r2_boot_full $BOOT > gpurun_out/c_full_boot.log 2>&1
note "boot full exit $?"
ls -la gpurun_out/*.ncu-rep 2>/dev/null; tail -n 2 gpurun_out/c_plain.log | cut -c1-600
