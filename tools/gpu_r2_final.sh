#!/bin/bash
# round-2 end evidence: the whole GPU suite, smoke, the default bench line and the reference arm on the same box, and the
# look-ahead compaction on a workload where 90 % of the prompts resolve at position 0
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
rm -f gpurun_out/summary_final.txt gpurun_out/full_depth_parity.jsonl gpurun_out/golden_position_mismatches.jsonl
run() { name=$1; shift; timeout ${TMO:-1500} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_final.txt; }
run h_tests   python -m pytest tests -x -q -m gpu --durations=6
run h_smoke   python __graft_entry__.py smoke
run h_bench   python bench.py
run h_ref     python bench.py --impl reference --steps 3 --warmup 1
B="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --no-other-modes --resolve-rate 0.9"
run h_res_compact $B
run h_res_full    $B --option lookahead_compact=0
tail -n 3 gpurun_out/h_tests.log; tail -n 1 gpurun_out/h_smoke.log; tail -n 1 gpurun_out/h_bench.log | cut -c1-3500; tail -n 1 gpurun_out/h_ref.log | cut -c1-1500
for f in h_res_compact h_res_full; do tail -n 1 gpurun_out/$f.log | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print('$f', round(d['value'], 1), 'prompts/s', d['config'].get('options'), d['gpu_launches'], 'launches')"; done
