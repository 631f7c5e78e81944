#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 1200 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -5 gpurun_out/pytest.log
python tools/ab_blend.py 2 6 0 > gpurun_out/ab_blend.json 2> gpurun_out/ab_blend.err; echo "ab rc=$?"; cat gpurun_out/ab_blend.json
python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
python bench.py --workload c1 --steps 10 --warmup 3 > gpurun_out/bench_c1_ours.json 2> gpurun_out/bench_c1_ours.err
python bench.py --workload slam --steps 10 --warmup 3 > gpurun_out/bench_slam_ours.json 2> gpurun_out/bench_slam_ours.err
