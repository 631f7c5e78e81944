#!/bin/bash
# the driver's round-end sequence on one GPU: tests, smoke, reference arm, our arm (default flags)
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/final_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/final_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/final_smoke.log
( time python bench.py --impl reference ) > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err; echo "ref rc=$?"
( time python bench.py ) > gpurun_out/final_bench_ours.json 2> gpurun_out/final_bench_ours.err; echo "ours rc=$?"
grep real gpurun_out/final_bench_reference.err gpurun_out/final_bench_ours.err
python tools/show_bench.py gpurun_out/final_bench_ours.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/final_bench_ours.json').read().strip().splitlines()[-1])
for k in ('value','e2e','roofline','roofline_hbm','cpu_baseline','drop_in_call_shape','train_step','ssim_ms','adam_ms','densify_ms','gpu_launches','clocks'):
    print(k, json.dumps(d.get(k))[:400])
r=json.loads(open('gpurun_out/final_bench_reference.json').read().strip().splitlines()[-1])
print('reference', r.get('value'), r.get('e2e'), r.get('cpu_baseline'))
PY
