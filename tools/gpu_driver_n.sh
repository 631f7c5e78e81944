#!/bin/bash
# the driver's N-GPU commands, exactly: reference arm first, then ours (default K/W)
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
( time $TR bench.py --impl reference --gpus $N --steps 5 --warmup 3 ) > gpurun_out/drv_ref_n$N.json 2> gpurun_out/drv_ref_n$N.err; echo "ref rc=$?"
( time $TR bench.py --gpus $N --steps 5 --warmup 3 ) > gpurun_out/drv_ours_n$N.json 2> gpurun_out/drv_ours_n$N.err; echo "ours rc=$?"
grep real gpurun_out/drv_ref_n$N.err gpurun_out/drv_ours_n$N.err
python - <<PY
import json
for f in ("ref","ours"):
    lines=[l for l in open("gpurun_out/drv_%s_n$N.json" % f).read().splitlines() if l.startswith("{")]
    d=json.loads(lines[-1]); print(f, len(lines), d.get("impl"), d.get("value"), d.get("n_gpus"), d.get("ms_per_step"), (d.get("e2e") or {}).get("value"), (d.get("parity_multi_gpu") or {}).get("ok"))
PY
