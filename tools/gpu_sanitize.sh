#!/bin/bash
# usage: bash tools/gpu_sanitize.sh memcheck|racecheck|synccheck|initcheck   (one tool per gpurun call, after a plain run)
TOOL=$1
mkdir -p gpurun_out
timeout 300 python tools/sanitize_case.py 3000 > gpurun_out/sanitize_plain.log 2>&1; rc=$?; echo "plain rc=$rc"; tail -2 gpurun_out/sanitize_plain.log
if [ $rc -ne 0 ]; then exit 86; fi
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 1 --print-limit 30 python tools/sanitize_case.py 3000 > gpurun_out/sanitize_$TOOL.log 2>&1; echo "$TOOL rc=$?"
tail -15 gpurun_out/sanitize_$TOOL.log
