#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --timeout 1200 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest.log
python bench.py --workload slam --steps 10 --warmup 3 > gpurun_out/bench_slam_ours.json 2>/dev/null
python bench.py --workload knn --steps 10 --warmup 3 > gpurun_out/bench_knn_ours.json 2>/dev/null
python bench.py --workload knn --gaussians 25000 --steps 50 --warmup 5 > gpurun_out/bench_knn25k_ours.json 2>/dev/null
python bench.py --impl reference --workload knn --gaussians 25000 --steps 50 --warmup 5 > gpurun_out/bench_knn25k_ref.json 2>/dev/null
python -c "
import json
for f in ('slam_ours','knn_ours','knn25k_ours','knn25k_ref'):
    d=json.loads(open('gpurun_out/bench_%s.json'%f).read().strip().splitlines()[-1]); print(f, d['value'], d.get('ms_per_step'), {k:v['ms_per_view'] for k,v in (d.get('kernels') or {}).items() if 'knn' in k})"
