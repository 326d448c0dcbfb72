#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary10.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary10.txt; }
nvidia-smi --query-gpu=name,power.limit,power.max_limit,clocks.max.sm,temperature.gpu --format=csv > gpurun_out/gpuinfo10.txt
for i in 1 2; do
run bench10_cg1_$i python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --gemm-cta-group 1
run bench10_cg2_$i python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --gemm-cta-group 2
done
cat gpurun_out/gpuinfo10.txt
for f in gpurun_out/bench10_*.log; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"],1), "prompts/s gemm", round(d["roofline"]["achieved"],1), "TF share", round(d["roofline"]["gemm_share_of_step"],3), "clk", d["clocks"]["sm_mhz"])
PY
done
