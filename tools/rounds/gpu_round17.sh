#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary17.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary17.txt; }
for cg in 0 1 2 3 0; do
run bench17_cg${cg}_$RANDOM python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --gemm-cta-group $cg
done
for f in gpurun_out/bench17_*.log; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], "cg", d["config"]["gemm_cta_group"], round(d["value"],1), "prompts/s gemm", round(d["roofline"]["achieved"],1), "TF share", round(d["roofline"]["gemm_share_of_step"],3), "clk", d["clocks"]["sm_mhz"])
PY
done
