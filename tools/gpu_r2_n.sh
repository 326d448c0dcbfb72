#!/bin/bash
# round 2, call N: ncu --set full for the kernels added late in the round (each capture follows a plain run of the same command that
# exited 0 in an earlier call of this round: tools/gpu_r2_k.sh for the Llama pass, tools/gpu_r2_t5c.sh for the T5 pass)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
L="python tools/family_bench.py --models llama2_7b --prompts 600"
T="python tools/t5_bench.py --prompts 600 --modes fp16 --steps 1"
timeout 600 $L > gpurun_out/n_plain_l.log 2>&1; echo "plain llama exit $?"
timeout 600 $T > gpurun_out/n_plain_t.log 2>&1; echo "plain t5 exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"attn_step_kernel|rope_cache_vec_kernel" -s 40 -c 4 -f -o gpurun_out/r2_step_full $L > gpurun_out/n_full_l.log 2>&1; echo "ncu llama exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"attn_enc_kernel|step_attn_kernel" -s 30 -c 6 -f -o gpurun_out/r2_t5_full $T > gpurun_out/n_full_t.log 2>&1; echo "ncu t5 exit $?"
ls -la gpurun_out/*.ncu-rep | tail -n 4
