#!/bin/bash
mkdir -p gpurun_out
cd "$(dirname "$0")/../.."
rm -f gpurun_out/summary33.txt
run() { name=$1; shift; timeout ${TMO:-900} "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary33.txt; }
run time33    python tools/traffic_sweep.py 131072 --time --pairs
run plain33   python tools/traffic_sweep.py 131072 --pairs
if grep -q "plain33 exit 0" gpurun_out/summary33.txt; then
run ncu33     ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,lts__t_sector_hit_rate.pct --clock-control none --csv --log-file gpurun_out/traffic33.csv -k regex:gemm_tc python tools/traffic_sweep.py 131072 --pairs
fi
grep '^{' gpurun_out/time33.log
