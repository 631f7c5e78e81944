#!/bin/bash
mkdir -p gpurun_out
python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref_c3.json 2> gpurun_out/bench_ref_c3.err; echo "c3 rc=$?"
for w in c1 c2 slam knn; do
  python bench.py --impl reference --workload $w --steps 10 --warmup 3 > gpurun_out/bench_ref_$w.json 2> gpurun_out/bench_ref_$w.err; echo "$w rc=$?"
done
python bench.py --workload c1 --steps 10 --warmup 3 > gpurun_out/bench_c1_ours.json 2>/dev/null
python bench.py --workload knn --steps 10 --warmup 3 > gpurun_out/bench_knn_ours.json 2>/dev/null
python bench.py --workload slam --aa --steps 10 --warmup 3 > gpurun_out/bench_slam_ours_aa.json 2>/dev/null
python -c "
import json
for f in ('ref_c3','ref_c1','ref_c2','ref_slam','ref_knn','c1_ours','knn_ours','slam_ours_aa'):
    try:
        d=json.loads(open('gpurun_out/bench_%s.json'%f).read().strip().splitlines()[-1]); print(f, d.get('value'), d.get('unit'), d.get('ms_per_step'))
    except Exception as e: print(f,'ERR',e)"
