#!/bin/bash
# round-end evidence: full GPU suite, smoke, default bench (both arms), ncu launch list + full capture of the top kernel
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_final.txt
run() { name=$1; shift; timeout ${TMO:-1500} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_final.txt; }
run f_tests   python -m pytest tests -x -q -m gpu
run f_smoke   python __graft_entry__.py smoke
run f_ref     python bench.py --impl reference --steps 3 --warmup 3
run f_bench   python bench.py
B="python bench.py --prompts 1200 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline"
run f_1k      $B
if grep -q "f_1k exit 0" gpurun_out/summary_final.txt; then
  run f_ncu_list ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_final.csv $B
  run f_ncu_full ncu --set full --clock-control none --import-source on -k regex:"gemm_tc_kernel" -s 2 -c 4 -f -o gpurun_out/gemm_full_final $B
fi
tail -n 3 gpurun_out/f_tests.log; tail -n 1 gpurun_out/f_smoke.log; tail -n 1 gpurun_out/f_ref.log | cut -c1-400; tail -n 1 gpurun_out/f_bench.log | cut -c1-2500
