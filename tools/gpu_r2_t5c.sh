#!/bin/bash
# round 2, call T5-c: where the time of the encoder-decoder pass goes (launch list of one small pass at T0_3B size), and the
# effect of the batch size on the one-token steps
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_t5c.txt
note() { echo "$*" | tee -a gpurun_out/summary_t5c.txt; }
timeout 600 python tools/t5_bench.py --prompts 600 --modes fp16 --steps 1 > gpurun_out/t5c_plain.jsonl 2> gpurun_out/t5c_plain.err; note "plain exit $?"
timeout 900 python tools/t5_bench.py --prompts 4000 --modes fp16 bf16_split --steps 1 --max-tokens 131072 > gpurun_out/t5c_big_batch.jsonl 2> gpurun_out/t5c_big.err; note "big batch exit $?"
if grep -q prompts_per_s gpurun_out/t5c_plain.jsonl; then
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/t5_launches.csv \
    python tools/t5_bench.py --prompts 600 --modes fp16 --steps 1 > gpurun_out/t5c_ncu.log 2>&1; note "ncu launch list exit $?"
fi
cat gpurun_out/t5c_plain.jsonl gpurun_out/t5c_big_batch.jsonl; wc -l gpurun_out/t5_launches.csv
