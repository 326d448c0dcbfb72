#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/ab38.jsonl
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "gemm_tc or lm_head" > gpurun_out/t_gemm38.log 2>&1; echo "t_gemm38 exit $?"; tail -n 2 gpurun_out/t_gemm38.log
for opt in "" "--option gemm_policy=0" "--option gemm_policy=0 --option gemm_wave_sync=0"; do
  timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e $opt 2>/dev/null | tail -n 1 >> gpurun_out/ab38.jsonl
done
python - <<'PY'
import json
for l in open('gpurun_out/ab38.jsonl'):
    j=json.loads(l); print(j['config'].get('options'), round(j['value'],1), round(j['roofline']['achieved'],1), j['clocks']['sm_mhz'])
PY
