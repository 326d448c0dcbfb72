#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
python tools/sweep_gemm.py 131072 1.5 up 0 > gpurun_out/sweep29_plain.log 2>&1
python tools/sweep_gemm.py 131072 1.5 up 1 > gpurun_out/sweep29_gelu.log 2>&1
python tools/sweep_gemm.py 131072 1.5 down 0 > gpurun_out/sweep29_down_plain.log 2>&1
python tools/sweep_gemm.py 131072 1.5 down 2 > gpurun_out/sweep29_down_resid.log 2>&1
for f in plain gelu down_plain down_resid; do echo "== $f"; grep '^{' gpurun_out/sweep29_$f.log; done
