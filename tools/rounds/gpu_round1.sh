#!/bin/bash
# first GPU pass: each group in its own process so that one faulting kernel does not hide the others
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
nproc > gpurun_out/host.txt; free -g >> gpurun_out/host.txt; nvidia-smi >> gpurun_out/host.txt 2>&1
run() { name=$1; shift; timeout 600 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary.txt; }
run t_gemm_f32   python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "gemm_f32"
run t_gemm_tc    python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "gemm_tc_store"
run t_gemm_epi   python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "epilogues or rejects"
run t_lmhead     python -m pytest tests/test_gpu_kernels.py -q -m gpu -k "lm_head"
run t_boot       python -m pytest tests/test_gpu_bootstrap.py -q -m gpu
run t_scorer32   python -m pytest tests/test_gpu_scorer.py -q -m gpu -k "fp32 or edge"
run t_scorerbf   python -m pytest tests/test_gpu_scorer.py -q -m gpu -k "bf16"
run gemm_bench   python tools/gemm_bench.py 32768
tail -n 5 gpurun_out/*.log
