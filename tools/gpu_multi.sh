#!/bin/bash
# usage: bash tools/gpu_multi.sh N [tests]   -- N-GPU measurements (torchrun, one rank per GPU); "tests" also runs the 2-rank pytest files
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
if [ "$2" == "tests" ]; then
  python -m pytest tests -m gpu -q --timeout 1500 -k "multirank or second_device" > gpurun_out/pytest_multi_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi_n$N.log
  tail -5 gpurun_out/pytest_multi_n$N.log
fi
$TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n${N}_weak.json 2> gpurun_out/bench_n${N}_weak.err; echo "weak rc=$?"
$TR bench.py --gpus $N --steps 10 --warmup 3 --scaling strong --no-e2e --no-parity > gpurun_out/bench_n${N}_strong.json 2> gpurun_out/bench_n${N}_strong.err; echo "strong rc=$?"
$TR bench.py --gpus $N --steps 10 --warmup 3 --workload train --no-e2e > gpurun_out/bench_n${N}_train.json 2> gpurun_out/bench_n${N}_train.err; echo "train rc=$?"
$TR tools/comm_check.py --json gpurun_out/comm_check_n$N.json > gpurun_out/comm_check_n$N.log 2>&1; echo "comm rc=$?"
python - <<PY
import json
for f in ("weak","strong","train"):
    try:
        d=json.load(open(f"gpurun_out/bench_n${N}_%s.json" % f)); print(f, d["value"], d["ms_per_step"], d.get("parity_multi_gpu"))
    except Exception as e: print(f, "ERR", e)
PY
