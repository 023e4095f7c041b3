#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for v in 2 3; do timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 $v 2 > gpurun_out/big_grid_v$v.log 2>&1; timeout 300 python tools/quick_perf.py stress 3840 2160 4 100 $v 2 > gpurun_out/big_stress_v$v.log 2>&1; done
timeout 300 python tools/quick_perf.py rtiow 1920 1080 64 50 3 2 > gpurun_out/big_rtiow_v3.log 2>&1
tail -n 3 gpurun_out/pytest_gpu.log gpurun_out/big_*.log
