#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/ab36.jsonl
for bt in 262144 280000 375000; do
  timeout 900 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --batch-tokens $bt 2>/dev/null | tail -n 1 >> gpurun_out/ab36.jsonl
done
python - <<'PY'
import json
for l in open('gpurun_out/ab36.jsonl'):
    j=json.loads(l); print(j['config']['batch_tokens'], j['config']['batches_per_rank'], round(j['value'],1), round(j['roofline']['achieved'],1), j['clocks']['sm_mhz'])
PY
