#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary28.txt gpurun_out/ab28.jsonl
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary28.txt; }
run t_gemm28 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "gemm_tc or lm_head"
for opt in "" "--option gemm_serpentine=0" ""; do
  timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e $opt 2>/dev/null | tail -n 1 >> gpurun_out/ab28.jsonl
done
MODELS="tiiuae/falcon-7b-instruct togethercomputer/RedPajama-INCITE-7B-Instruct"
run c4_a python -m llm_interpretation_replication_b200.entrypoints.compare_instruct_models --layers 2 --questions 24 --models $MODELS --out gpurun_out/c4_a
B2S_MAX_TOKENS=700 run c4_b python -m llm_interpretation_replication_b200.entrypoints.compare_instruct_models --layers 2 --questions 11 --models $MODELS --out gpurun_out/c4_b
tail -n 3 gpurun_out/t_gemm28.log
python - <<'PY'
import json, pandas as pd
for l in open('gpurun_out/ab28.jsonl'):
    j=json.loads(l); print(j['config'].get('options'), round(j['value'],1), round(j['roofline']['achieved'],1), j['clocks']['sm_mhz'])
a=pd.read_csv("gpurun_out/c4_a/instruct_model_comparison_results.csv"); b=pd.read_csv("gpurun_out/c4_b/instruct_model_comparison_results.csv")
m=a.merge(b,on=["prompt","model"],suffixes=("_a","_b"))
print("rows in both", len(m), "max |d rel|", (m.relative_prob_a-m.relative_prob_b).abs().max(), "outputs differing", (m.model_output_a.fillna('')!=m.model_output_b.fillna('')).sum())
PY
