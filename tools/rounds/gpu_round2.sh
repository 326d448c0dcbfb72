#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary2.txt; }
run t_boot2      python -m pytest tests/test_gpu_bootstrap.py -q -m gpu
run t_scorer2    python -m pytest tests/test_gpu_scorer.py -q -m gpu -k "edge or real_dims"
run smoke        python __graft_entry__.py smoke
run bench_quick  python bench.py --steps 1 --warmup 1 --no-cpu-baseline
run bench_1k     python bench.py --prompts 1000 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline
if grep -q "bench_1k exit 0" gpurun_out/summary2.txt; then
  TMO=1200 run ncu_list ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_r1.csv python bench.py --prompts 1000 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline
fi
TMO=600 run ref_arm python bench.py --impl reference --steps 1 --warmup 0
tail -n 4 gpurun_out/t_boot2.log gpurun_out/t_scorer2.log gpurun_out/smoke.log gpurun_out/bench_quick.log gpurun_out/bench_1k.log gpurun_out/ref_arm.log
