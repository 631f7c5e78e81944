#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --timeout 1200 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest.log
python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench rc=$?"
GSR_SCAN_TWO_PASS=0 python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline --no-e2e > gpurun_out/bench_chained.json 2> gpurun_out/bench_chained.err; echo "bench chained rc=$?"
python tools/show_bench.py gpurun_out/bench_ours.json gpurun_out/bench_chained.json
python bench.py --workload slam --steps 10 --warmup 3 > gpurun_out/bench_slam_ours.json 2>/dev/null
python bench.py --workload c2 --steps 10 --warmup 3 > gpurun_out/bench_c2_ours.json 2>/dev/null
python -c "
import json
for f in ('slam','c2'):
    d=json.loads(open('gpurun_out/bench_%s_ours.json'%f).read().strip().splitlines()[-1]); print(f, d['value'], d['ms_per_view'])"
