#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --timeout 1200 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest.log
for v in 1 0 1 0; do
  GSR_SPLIT_PRE=$v python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline --no-e2e > gpurun_out/bench_split$v.json 2> gpurun_out/bench_split$v.err; echo "split=$v rc=$?"
  python tools/show_bench.py gpurun_out/bench_split$v.json | head -16
done
for v in 1 0; do
  GSR_SPLIT_PRE=$v python bench.py --workload slam --steps 10 --warmup 3 > gpurun_out/bench_slam_split$v.json 2>/dev/null
  GSR_SPLIT_PRE=$v python bench.py --workload c2 --steps 10 --warmup 3 > gpurun_out/bench_c2_split$v.json 2>/dev/null
done
python -c "
import json
for f in ('slam_split1','slam_split0','c2_split1','c2_split0'):
    d=json.loads(open('gpurun_out/bench_%s.json'%f).read().strip().splitlines()[-1]); print(f, d['value'], d.get('ms_per_view'))"
