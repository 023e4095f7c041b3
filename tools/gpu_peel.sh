#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for np in 0 1; do
  if [ $np = 1 ]; then export CRT_NO_PEEL=1; else unset CRT_NO_PEEL; fi
  for v in 1 3; do timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 $v 2 > gpurun_out/peel${np}_rtiow_v$v.log 2>&1; done
  timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 3 2 > gpurun_out/peel${np}_grid_v3.log 2>&1
  timeout 300 python tools/quick_perf.py stress 3840 2160 4 100 3 2 > gpurun_out/peel${np}_stress_v3.log 2>&1
  timeout 300 python tools/quick_perf.py textured 1920 1080 16 50 1 2 > gpurun_out/peel${np}_textured_v1.log 2>&1
  timeout 300 python tools/quick_perf.py default 1200 800 64 50 1 2 > gpurun_out/peel${np}_default_v1.log 2>&1
done
tail -n 3 gpurun_out/pytest_gpu.log; for f in gpurun_out/peel*.log; do echo $f; grep -h "Mrays" $f | tail -1; done
