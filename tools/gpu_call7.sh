#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline --no-e2e > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
GSR_LIB=$PWD/gaussian-splatting-slam_b200/libgsr_sort2048.so python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline --no-e2e > gpurun_out/bench_sort2048.json 2> gpurun_out/bench_sort2048.err; echo "bench sort2048 rc=$?"
python -m pytest tests -m gpu -q --timeout 1200 -k "c1 or ragged or grad_sinks or binning or c2" > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -3 gpurun_out/pytest.log
python tools/prof_view.py 4 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"scan_offsets|onesweep_kernel" -s 12 -c 7 -o gpurun_out/prof_binning python tools/prof_view.py 4 > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"
