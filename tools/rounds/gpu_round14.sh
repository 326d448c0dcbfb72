#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary14.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary14.txt; }
run t_attn14   python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "attention"
run cmp_plain  python tools/ncu_gemm_compare.py 32768
if grep -q "cmp_plain exit 0" gpurun_out/summary14.txt; then
  run ncu_cmp ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,smsp__inst_executed.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__cycles_elapsed.avg.per_second,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,smsp__inst_executed_op_shared_ld.sum --clock-control none -k regex:"gemm|nvjet|cutlass|xmma|sm100|cublas" --csv --log-file gpurun_out/ncu_cmp.csv python tools/ncu_gemm_compare.py 32768
fi
tail -n 4 gpurun_out/t_attn14.log
