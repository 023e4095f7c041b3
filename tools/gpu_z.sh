#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for v in 1 2 3; do timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 $v > gpurun_out/z_rtiow_v$v.log 2>&1; done
for sc in textured stress grid; do for v in 1 3; do timeout 300 python tools/quick_perf.py $sc 1920 1080 16 50 $v > gpurun_out/z_${sc}_v$v.log 2>&1; done; done
timeout 120 python tools/quick_perf.py default 1200 800 64 50 1 > gpurun_out/z_default_v1.log 2>&1
tail -n 4 gpurun_out/z_*.log
