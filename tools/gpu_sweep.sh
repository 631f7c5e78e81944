#!/bin/bash
N=$1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 tools/comm_sweep.py --json gpurun_out/comm_sweep_n$N.json > gpurun_out/comm_sweep_n$N.log 2>&1; echo "sweep rc=$?"
grep '^{"variant"' gpurun_out/comm_sweep_n$N.log
tail -5 gpurun_out/comm_sweep_n$N.log | cut -c1-300
