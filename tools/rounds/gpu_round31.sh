#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/ab31.jsonl
for bt in 131072 262144 393216 589824; do
  timeout 900 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --batch-tokens $bt 2>gpurun_out/ab31_$bt.err | tail -n 1 >> gpurun_out/ab31.jsonl
done
python - <<'PY'
import json
for l in open('gpurun_out/ab31.jsonl'):
    try:
        j=json.loads(l); print(j['config']['batch_tokens'], j['config']['batches_per_rank'], round(j['value'],1), round(j['roofline']['achieved'],1), j['clocks']['sm_mhz'], j['gpu_launches'])
    except Exception as e: print('bad line', l[:200])
PY
tail -n 3 gpurun_out/ab31_*.err | cut -c1-300
