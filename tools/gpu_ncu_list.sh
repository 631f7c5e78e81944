#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-e2e --no-kernel-timing"
$CMD > gpurun_out/ncu_plain_run.json 2> gpurun_out/ncu_plain_run.err; rc=$?; echo "plain rc=$rc"
if [ $rc -ne 0 ]; then exit 86; fi
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/ncu_launches.csv $CMD > gpurun_out/ncu_list_run.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/ncu_launches.csv
