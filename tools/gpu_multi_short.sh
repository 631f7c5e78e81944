#!/bin/bash
# usage: bash tools/gpu_multi_short.sh N   -- weak + strong lines only (see tools/gpu_multi.sh for the full set)
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
$TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n${N}_weak.json 2> gpurun_out/bench_n${N}_weak.err; echo "weak rc=$?"
$TR bench.py --gpus $N --steps 10 --warmup 3 --scaling strong --no-e2e --no-parity > gpurun_out/bench_n${N}_strong.json 2> gpurun_out/bench_n${N}_strong.err; echo "strong rc=$?"
python - <<PY
import json
for f in ("weak","strong"):
    d=json.load(open("gpurun_out/bench_n${N}_%s.json" % f)); print(f, d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value"), (d.get("parity_multi_gpu") or {}).get("max"), (d.get("parity_multi_gpu") or {}).get("ok"))
PY
