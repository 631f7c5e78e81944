#!/bin/bash
mkdir -p gpurun_out
GSR_BINDING=ctypes python tools/ab_blend.py 2 > gpurun_out/ab_pf1.json 2> gpurun_out/ab_pf1.err; echo "pf1 rc=$?"; cat gpurun_out/ab_pf1.json
GSR_LIB=$PWD/gaussian-splatting-slam_b200/libgsr_nopf.so python tools/ab_blend.py 2 > gpurun_out/ab_pf0.json 2> gpurun_out/ab_pf0.err; echo "pf0 rc=$?"; cat gpurun_out/ab_pf0.json
GSR_BINDING=ctypes python tools/ab_blend.py 2 > gpurun_out/ab_pf1b.json 2> gpurun_out/ab_pf1b.err; cat gpurun_out/ab_pf1b.json
