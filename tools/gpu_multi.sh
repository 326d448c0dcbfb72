#!/bin/bash
# 2-GPU checks: bench at N=2 (NCCL all-gather of the rows) and the sharded entry point 2 (BASELINE config 4)
# against the same entry point on one GPU (tables must be identical: weights replicated, per-prompt math independent).
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_multi.txt
N=${1:-2}
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_multi.txt; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run multi_bench   $TR --master-port 29551 bench.py --gpus $N --steps 2 --warmup 1
run multi_ref     $TR --master-port 29552 bench.py --impl reference --gpus $N --steps 1 --warmup 0
MODELS="tiiuae/falcon-7b-instruct togethercomputer/RedPajama-INCITE-7B-Instruct"
run c4_sharded    $TR --master-port 29553 -m llm_interpretation_replication_b200.entrypoints.compare_instruct_models --layers 2 --questions 24 --models $MODELS --out gpurun_out/c4_n$N
run c4_single     python -m llm_interpretation_replication_b200.entrypoints.compare_instruct_models --layers 2 --questions 24 --models $MODELS --out gpurun_out/c4_n1
python - <<PY | tee -a gpurun_out/summary_multi.txt
import pandas as pd
a = pd.read_csv("gpurun_out/c4_n1/instruct_model_comparison_results.csv")
b = pd.read_csv("gpurun_out/c4_n$N/instruct_model_comparison_results.csv")
same = a.equals(b)
print("config-4 tables identical at 1 and $N ranks:", same, len(a), "rows")
PY
tail -n 3 gpurun_out/multi_bench.log | cut -c1-1200; tail -n 2 gpurun_out/multi_ref.log | cut -c1-600
