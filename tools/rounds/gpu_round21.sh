#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary21.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary21.txt; }
run attn21  python tools/attn_bench.py
if grep -q "attn21 exit 0" gpurun_out/summary21.txt; then
run ncu21   ncu --set full --clock-control none --import-source on -k regex:attn_tc --launch-skip 2 -c 1 -o gpurun_out/attn_full_r1 -f python tools/attn_bench.py
fi
cat gpurun_out/attn21.log
