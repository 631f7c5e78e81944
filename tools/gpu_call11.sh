#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/noalloc_alone.log
python -m pytest tests -m gpu -x -q --timeout 1200 > gpurun_out/pytest.log 2>&1; echo "all rc=$?"; tail -3 gpurun_out/pytest.log
