#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary32.txt
run() { name=$1; shift; timeout ${TMO:-600} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary32.txt; }
T="python -m pytest tests/test_gpu_kernels.py tests/test_gpu_corr.py -x -q -m gpu -k attention_bf16-tc_p_in_tmem or simt_kv or corr_kernel or rejects"
run plain32 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_corr.py -x -q -m gpu -k "tc_p_in_tmem or simt_kv or corr_kernel or rejects"
if grep -q "plain32 exit 0" gpurun_out/summary32.txt; then
run memcheck32 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_corr.py -x -q -m gpu -k "tc_p_in_tmem or simt_kv or corr_kernel or rejects"
fi
tail -n 3 gpurun_out/plain32.log; grep -E "ERROR SUMMARY|passed|failed|Invalid|Error" gpurun_out/memcheck32.log | head -12
