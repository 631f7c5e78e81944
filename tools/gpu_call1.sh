#!/bin/bash
# GPU call: tests + all bench workloads (ours and reference arm)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 1200 -x > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -15 gpurun_out/pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
for wl in c1 c2 slam knn; do
  for impl in ours reference; do
    python bench.py --workload $wl --impl $impl --steps 10 --warmup 3 > gpurun_out/bench_${wl}_${impl}.json 2> gpurun_out/bench_${wl}_${impl}.err; echo "$wl $impl rc=$?"
  done
done
python bench.py --workload slam --aa --steps 10 --warmup 3 > gpurun_out/bench_slam_aa_ours.json 2> gpurun_out/bench_slam_aa.err
python bench.py --workload c1 --aa --steps 10 --warmup 3 > gpurun_out/bench_c1_aa_ours.json 2> gpurun_out/bench_c1_aa.err
tail -c 600 gpurun_out/bench_ours.err
head -c 1500 gpurun_out/bench_ours.json
