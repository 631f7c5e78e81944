#!/bin/bash
mkdir -p gpurun_out
T=tests/test_gpu_binning.py::test_num_rendered_beyond_2_pow_30_is_refused_on_both_routes
SEGV_TRACE_FILE=$PWD/gpurun_out/segv_trace.txt LD_PRELOAD=$PWD/tools/debug/segv_trace.so timeout 300 python -m pytest $T -q -x -s -p no:faulthandler > gpurun_out/t30_ext.log 2>&1; echo "ext rc=$?"
head -40 gpurun_out/segv_trace.txt
python bench.py --workload slam --steps 10 --warmup 3 > gpurun_out/slam_a.json 2>/dev/null
python bench.py --workload slam --steps 10 --warmup 10 > gpurun_out/slam_b.json 2>/dev/null
GSR_BENCH_NO_SAMPLER=1 python bench.py --workload slam --steps 10 --warmup 3 > gpurun_out/slam_c.json 2>/dev/null
python bench.py --workload slam --steps 10 --warmup 3 --no-kernel-timing > gpurun_out/slam_d.json 2>/dev/null
python bench.py --workload slam --steps 40 --warmup 3 --no-kernel-timing > gpurun_out/slam_e.json 2>/dev/null
for f in a b c d e; do python -c "
import json;d=json.loads(open('gpurun_out/slam_$f.json').read().strip().splitlines()[-1]);print('$f',d['value'],d['ms_per_view'],d['cuda_mallocs_in_timed_region'],d['clocks'].get('samples'))"; done
