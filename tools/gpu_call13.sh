#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --timeout 1200 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest.log
for v in 1 0; do
  GSR_SMALL_SORT=$v python bench.py --workload c1 --steps 20 --warmup 5 > gpurun_out/bench_c1_ss$v.json 2>/dev/null
  GSR_SMALL_SORT=$v python bench.py --workload c1 --steps 20 --warmup 5 > gpurun_out/bench_c1_ss${v}b.json 2>/dev/null
done
python bench.py --workload slam --steps 10 --warmup 3 > gpurun_out/bench_slam_ours.json 2>/dev/null
python bench.py --impl reference --workload c1 --steps 20 --warmup 5 > gpurun_out/bench_ref_c1.json 2>/dev/null
python -c "
import json
for f in ('c1_ss1','c1_ss1b','c1_ss0','c1_ss0b','slam_ours','ref_c1'):
    d=json.loads(open('gpurun_out/bench_%s.json'%f).read().strip().splitlines()[-1]); print(f, d['value'], d.get('ms_per_view'), {k:v['ms_per_view'] for k,v in (d.get('kernels') or {}).items() if 'sort' in k or 'onesweep' in k or 'hist' in k})"
