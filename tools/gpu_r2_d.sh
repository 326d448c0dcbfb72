#!/bin/bash
# round 2, call D: new bootstrap kernels and GEMM knobs -- tests, bootstrap bench, same-box end-to-end A/B of the knobs
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_d.txt gpurun_out/d_ab.jsonl
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_d.txt; }
run d_tests python -m pytest tests/test_gpu_bootstrap.py tests/test_gpu_kernels.py tests/test_gpu_corr.py tests/test_gpu_agreement.py tests/test_gpu_entrypoints.py -q --maxfail=10 -rf
run d_boot python tools/boot_bench.py
B="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-other-modes"
for cfg in "" "--option gemm_prefetch=8" "--option gemm_band_n=-1" "--option gemm_prefetch=8 --option gemm_band_n=-1" "--option gemm_prefetch=16" "--option gemm_prefetch=4" "" "--option gemm_prefetch=8"; do
  timeout 400 $B $cfg 2> gpurun_out/d_ab_err.log | tail -n 1 | python -c "
import sys, json
l = sys.stdin.readline()
try:
    d = json.loads(l)
    print(json.dumps({'options': d['config']['options'], 'value': round(d['value'], 1), 'gemm_tflops': round(d['roofline']['achieved'], 1), 'frac': round(d['roofline']['frac'], 4), 'sm_mhz': d['clocks'].get('sm_mhz'), 'finite': d['config']['all_rel_prob_finite']}))
except Exception as e:
    print(json.dumps({'error': str(e), 'line': l[:200]}))
" | tee -a gpurun_out/d_ab.jsonl
done
tail -n 6 gpurun_out/d_tests.log; tail -n 1 gpurun_out/d_boot.log
