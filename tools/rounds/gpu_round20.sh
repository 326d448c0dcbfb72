#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary20.txt
run() { name=$1; shift; timeout ${TMO:-1200} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary20.txt; }
run t_serp20  python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "serpentine"
if grep -q "t_serp20 exit 0" gpurun_out/summary20.txt; then
run time20    python tools/traffic_sweep.py 111104 --time
run plain20   python tools/traffic_sweep.py 111104
if grep -q "plain20 exit 0" gpurun_out/summary20.txt; then
run ncu20     ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none --csv --log-file gpurun_out/traffic20.csv -k regex:gemm_tc python tools/traffic_sweep.py 111104
fi
fi
tail -n 3 gpurun_out/t_serp20.log; cat gpurun_out/time20.log
