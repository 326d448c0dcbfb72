#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary35.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary35.txt; }
run attn35  python tools/attn_bench.py typical
if grep -q "attn35 exit 0" gpurun_out/summary35.txt; then
run ncu35   ncu --set full --clock-control none --import-source on -k regex:attn_tc2 --launch-skip 8 -c 1 -o gpurun_out/attn2_full_r1 -f python tools/attn_bench.py typical
fi
cat gpurun_out/attn35.log
