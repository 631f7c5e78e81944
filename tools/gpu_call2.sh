#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 1200 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -30 gpurun_out/pytest.log
python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench ours rc=$?"
for wl in slam knn; do
    python bench.py --workload $wl --steps 10 --warmup 3 > gpurun_out/bench_${wl}_ours.json 2> gpurun_out/bench_${wl}_ours.err; echo "$wl rc=$?"
done
python tools/diag_host.py c1 ours > gpurun_out/diag_host_c1_ours.txt 2>&1
python tools/diag_host.py c1 reference > gpurun_out/diag_host_c1_ref.txt 2>&1
head -c 600 gpurun_out/diag_host_c1_ours.txt
