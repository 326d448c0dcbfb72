#!/bin/bash
# round 2, scaling call: bench.py at N GPUs of one box under torchrun (prompt-sharded, replicated weights, NCCL all-gather of the rows)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561"
timeout 1200 $TR bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/scale_n$N.log 2> gpurun_out/scale_n$N.err; echo "bench N=$N exit $?" | tee gpurun_out/summary_scale_n$N.txt
grep '^{' gpurun_out/scale_n$N.log | tail -n 1 | cut -c1-1800; tail -n 3 gpurun_out/scale_n$N.err
