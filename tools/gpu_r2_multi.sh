#!/bin/bash
# round 2, 2-GPU call: two devices in one process, bench at N = 2 (NCCL all-gather of the rows), sharded entry points against
# the single-GPU tables (must be identical: weights replicated, a prompt's row does not depend on its batch or step)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_multi.txt
N=${1:-2}
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_multi.txt; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run m_twodev      python -m pytest tests/test_gpu_scorer.py -q -k "two_devices" -rs
run m_bench       $TR --master-port 29551 bench.py --gpus $N --steps 2 --warmup 1
MODELS="tiiuae/falcon-7b-instruct togethercomputer/RedPajama-INCITE-7B-Instruct bigscience/T0_3B Qwen/Qwen-7B-Chat"
run m_c4_sharded  $TR --master-port 29553 -m llm_interpretation_replication_b200.entrypoints.compare_instruct_models --layers 2 --questions 24 --models $MODELS --out gpurun_out/c4_n$N
run m_c4_single   python -m llm_interpretation_replication_b200.entrypoints.compare_instruct_models --layers 2 --questions 24 --models $MODELS --out gpurun_out/c4_n1
run m_p_sharded   $TR --master-port 29554 -m llm_interpretation_replication_b200.entrypoints.score_perturbations --layers 2 --limit 300 --out gpurun_out/pert_n$N
run m_p_single    python -m llm_interpretation_replication_b200.entrypoints.score_perturbations --layers 2 --limit 300 --out gpurun_out/pert_n1
python - <<PY | tee -a gpurun_out/summary_multi.txt
import pandas as pd
a = pd.read_csv("gpurun_out/c4_n1/instruct_model_comparison_results.csv")
b = pd.read_csv("gpurun_out/c4_n$N/instruct_model_comparison_results.csv")
print("config-4 tables identical at 1 and $N ranks:", a.equals(b), len(a), "rows")
a = pd.read_csv("gpurun_out/pert_n1/perturbation_results_local.csv")
b = pd.read_csv("gpurun_out/pert_n$N/perturbation_results_local.csv")
print("perturbation tables identical at 1 and $N ranks:", a.equals(b), len(a), "rows")
PY
tail -n 3 gpurun_out/m_twodev.log; tail -n 1 gpurun_out/m_bench.log | cut -c1-1500
