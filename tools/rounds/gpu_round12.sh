#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary12.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary12.txt; }
run t_f16_gemm   python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "fp16_operands"
run t_scorer12   python -m pytest tests/test_gpu_scorer.py -q -m gpu
run parity12     python tools/full_size_parity.py 24 32
run bench12_bf   python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --mode bf16
run bench12_f16  python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --mode fp16
run t_ci         python -m pytest tests/test_gpu_bootstrap.py -q -m gpu -k 'ci_ or entry_point_4'
tail -n 6 gpurun_out/t_f16_gemm.log gpurun_out/t_scorer12.log; tail -n 4 gpurun_out/parity12.log; tail -n 15 gpurun_out/t_ci.log
for f in gpurun_out/bench12_*.log; do python - "$f" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"],1), "prompts/s gemm", round(d["roofline"]["achieved"],1), "TF share", round(d["roofline"]["gemm_share_of_step"],3), "clk", d["clocks"]["sm_mhz"])
PY
done
