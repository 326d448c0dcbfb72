#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary19.txt
run() { name=$1; shift; timeout ${TMO:-1500} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary19.txt; }
run diag19   python tools/diag_neox_bf16.py
run t_all19  python -m pytest tests -q -m gpu
cat gpurun_out/diag19.log | cut -c1-600; tail -n 8 gpurun_out/t_all19.log
