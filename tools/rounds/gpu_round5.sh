#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary5.txt
run() { name=$1; shift; timeout ${TMO:-1200} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary5.txt; }
run t_noalloc_alone python -m pytest tests/test_zz_no_alloc.py -q -m gpu
run t_all_gpu       python -m pytest tests -x -q -m gpu
tail -n 6 gpurun_out/t_noalloc_alone.log gpurun_out/t_all_gpu.log
