#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 1200 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -5 gpurun_out/pytest.log
python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
python bench.py --workload c1 --steps 10 --warmup 3 > gpurun_out/bench_c1_ours.json 2> gpurun_out/bench_c1_ours.err
python bench.py --workload slam --steps 10 --warmup 3 > gpurun_out/bench_slam_ours.json 2> gpurun_out/bench_slam_ours.err
python bench.py --workload knn --steps 10 --warmup 3 > gpurun_out/bench_knn_ours.json 2> gpurun_out/bench_knn_ours.err
python tools/diag_host.py c1 ours > gpurun_out/diag_host_c1_ext.txt 2>&1
GSR_BINDING=ctypes python tools/diag_host.py c1 ours > gpurun_out/diag_host_c1_ctypes.txt 2>&1
GSR_BINDING=ctypes python bench.py --workload c1 --steps 10 --warmup 3 > gpurun_out/bench_c1_ctypes.json 2> gpurun_out/bench_c1_ctypes.err
head -1 gpurun_out/diag_host_c1_ext.txt gpurun_out/diag_host_c1_ctypes.txt
