#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 1200 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -5 gpurun_out/pytest.log
python tools/ab_blend.py 0 1 > gpurun_out/ab_blend.json 2> gpurun_out/ab_blend.err; echo "ab rc=$?"; cat gpurun_out/ab_blend.json
python tools/prof_view.py 4 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:scan_emit -s 2 -c 1 -o gpurun_out/prof_scan_emit python tools/prof_view.py 4 > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"
