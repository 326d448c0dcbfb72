#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/ab25.jsonl
for rep in 1 2; do
for opt in "" "--option gemm_small_w_group=0" "--option gemm_serpentine=0" "--option attn_ts=0"; do
  timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e $opt 2>/dev/null | tail -n 1 >> gpurun_out/ab25.jsonl
done
done
python - <<'PY'
import json
for l in open('gpurun_out/ab25.jsonl'):
    j=json.loads(l); print(j['config'].get('options'), round(j['value'],1), round(j['roofline']['achieved'],1), j['clocks']['sm_mhz'])
PY
