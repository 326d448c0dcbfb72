#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary23.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary23.txt; }
run time23    python tools/traffic_sweep.py 111104 --time --hints
run plain23   python tools/traffic_sweep.py 111104 --hints
if grep -q "plain23 exit 0" gpurun_out/summary23.txt; then
run ncu23     ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct --clock-control none --csv --log-file gpurun_out/traffic23.csv -k regex:gemm_tc python tools/traffic_sweep.py 111104 --hints
fi
cat gpurun_out/time23.log
