#!/bin/bash
# round 2, call H: the whole GPU suite after the encoder-decoder branch and the q|k|v-bias Llama block went in, smoke, T5 evidence
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_h.txt gpurun_out/t5_golden_parity.jsonl gpurun_out/full_depth_parity.jsonl gpurun_out/golden_position_mismatches.jsonl
note() { echo "$*" | tee -a gpurun_out/summary_h.txt; }
timeout 2400 python -m pytest tests -q -m gpu --maxfail=15 -rf --durations=8 > gpurun_out/hh_tests.log 2>&1; note "gpu suite exit $?"
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/hh_smoke.log 2>&1; note "smoke exit $?"
tail -n 25 gpurun_out/hh_tests.log | cut -c1-250; tail -n 2 gpurun_out/hh_smoke.log; cat gpurun_out/t5_golden_parity.jsonl | tail -n 8
