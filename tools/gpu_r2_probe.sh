#!/bin/bash
# round 2: where does the tensor pipe idle?  One GEMM launch per (layer shape, knob) under an ncu metrics pass.
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
python tools/gemm_probe.py > gpurun_out/probe_plain.jsonl 2> gpurun_out/probe_plain.err &&
ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,sm__cycles_elapsed.avg.per_second \
    --clock-control none -k regex:gemm_tc --csv --log-file gpurun_out/gemm_probe.csv python tools/gemm_probe.py > gpurun_out/probe_ncu.jsonl 2> gpurun_out/probe_ncu.err
echo "probe exit $?"
wc -l gpurun_out/probe_ncu.jsonl gpurun_out/gemm_probe.csv
