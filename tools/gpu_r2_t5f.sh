#!/bin/bash
# round 2, call T5-f: launch list of the encoder-decoder pass at the default batch size (131072 tokens), T0_3B size, FP16
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python tools/t5_bench.py --prompts 2400 --modes fp16 --steps 1 --max-tokens 131072 --max-seqs 2048 > gpurun_out/t5f_plain.jsonl 2> gpurun_out/t5f_plain.err; echo "plain exit $?"
if grep -q prompts_per_s gpurun_out/t5f_plain.jsonl; then
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/t5_launches_big.csv \
    python tools/t5_bench.py --prompts 2400 --modes fp16 --steps 1 --max-tokens 131072 --max-seqs 2048 > gpurun_out/t5f_ncu.log 2>&1; echo "ncu exit $?"
fi
cat gpurun_out/t5f_plain.jsonl | cut -c1-300; wc -l gpurun_out/t5_launches_big.csv
