#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/t_all34.log 2>&1; echo "t_all34 exit $?"
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/t_all34.log | head -30
