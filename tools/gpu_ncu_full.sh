#!/bin/bash
# one `ncu --set full` run over the kernels of ONE view of the C3 bench (after the same command ran clean without ncu)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-extras --no-cpu-baseline --no-e2e --no-kernel-timing --no-parity"
$CMD > gpurun_out/ncu_plain_run.json 2> gpurun_out/ncu_plain_run.err; rc=$?; echo "plain rc=$rc"
if [ $rc -ne 0 ]; then exit 86; fi
# skip the warm-up (3 steps x 8 views x ~14 launches) and take the next 16 launches of the library's kernels
ncu --set full --clock-control none --import-source on -k regex:'blend|preprocess|onesweep|emit_inst|scan_|tile_table|radix_hist' --launch-skip 375 -c 15 -f -o gpurun_out/r02_full $CMD > gpurun_out/ncu_full_run.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/r02_full.ncu-rep
