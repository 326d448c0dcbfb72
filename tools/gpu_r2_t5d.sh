#!/bin/bash
# round 2, call T5-d: batch size sweep of the encoder-decoder pass at T0_3B size, and the launch list of the current kernels
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_t5d.txt gpurun_out/t5d_sweep.jsonl
note() { echo "$*" | tee -a gpurun_out/summary_t5d.txt; }
for mt in 32768 131072 262144; do
  timeout 900 python tools/t5_bench.py --prompts 4800 --modes fp16 --steps 1 --max-tokens $mt --max-seqs 4096 >> gpurun_out/t5d_sweep.jsonl 2> gpurun_out/t5d_sweep.err; note "sweep $mt exit $?"
done
timeout 900 python tools/t5_bench.py --prompts 4800 --modes bf16_split --steps 1 --max-tokens 262144 --max-seqs 4096 >> gpurun_out/t5d_sweep.jsonl 2>> gpurun_out/t5d_sweep.err; note "split exit $?"
timeout 600 python tools/t5_bench.py --prompts 600 --modes fp16 --steps 1 > gpurun_out/t5d_plain.jsonl 2> gpurun_out/t5d_plain.err; note "plain exit $?"
if grep -q prompts_per_s gpurun_out/t5d_plain.jsonl; then
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/t5_launches_after.csv \
    python tools/t5_bench.py --prompts 600 --modes fp16 --steps 1 > gpurun_out/t5d_ncu.log 2>&1; note "ncu launch list exit $?"
fi
cat gpurun_out/t5d_sweep.jsonl; wc -l gpurun_out/t5_launches_after.csv; tail -3 gpurun_out/t5d_sweep.err
