#!/bin/bash
mkdir -p gpurun_out
T=tests/test_gpu_binning.py::test_num_rendered_beyond_2_pow_30_is_refused_on_both_routes
GSR_BINDING=ctypes timeout 300 python -m pytest $T -q -x > gpurun_out/t30_ctypes.log 2>&1; echo "ctypes rc=$?"
LD_PRELOAD=$PWD/tools/debug/segv_trace.so timeout 300 python -m pytest $T -q -x -p no:faulthandler > gpurun_out/t30_ext.log 2>&1; echo "ext rc=$?"
tail -3 gpurun_out/t30_ctypes.log; grep -n "backtrace" -A 30 gpurun_out/t30_ext.log | head -50
timeout 1200 python -m pytest tests -m gpu -q --timeout 1200 --deselect $T > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -5 gpurun_out/pytest.log
