#!/bin/bash
# round 2, call M: streaming one-token attention through a cp.async ring: parity tests, then the Llama-2-7B launch list again
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests/test_gpu_scorer.py -q --maxfail=10 -rf > gpurun_out/m_scorer.log 2>&1; echo "scorer tests exit $?"
timeout 600 python -m pytest tests/test_gpu_full_depth.py -q -k "c1 or c2" -rf > gpurun_out/m_depth.log 2>&1; echo "full depth C1 C2 exit $?"
tail -n 4 gpurun_out/m_scorer.log | cut -c1-250; tail -n 3 gpurun_out/m_depth.log | cut -c1-250
bash tools/gpu_r2_k.sh
