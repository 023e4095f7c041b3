#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lbvh.py -m gpu -x -q > gpurun_out/pytest_lbvh.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_lbvh.log
for b in host gpu; do
  CRT_BVH_BUILDER=$b timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 3 2 > gpurun_out/lbvh_grid_$b.log 2>&1
  CRT_BVH_BUILDER=$b timeout 300 python tools/quick_perf.py stress 3840 2160 4 100 3 2 > gpurun_out/lbvh_stress_$b.log 2>&1
  CRT_BVH_BUILDER=$b timeout 300 python tools/quick_perf.py rtiow 1920 1080 64 50 3 2 > gpurun_out/lbvh_rtiow_$b.log 2>&1
done
tail -n 30 gpurun_out/pytest_lbvh.log; grep -h "scene\|Mrays" gpurun_out/lbvh_*.log
