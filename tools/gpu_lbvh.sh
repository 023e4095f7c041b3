#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lbvh.py tests/test_gpu_edit.py -m gpu -x -q -s > gpurun_out/pytest_lbvh.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_lbvh.log
for b in host lbvh ploc; do
  CRT_BVH_TIMING=1 CRT_BVH_BUILDER=$b timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 3 2 > gpurun_out/lbvh_grid_$b.log 2>&1
  CRT_BVH_BUILDER=$b timeout 300 python tools/quick_perf.py stress 3840 2160 4 100 3 2 > gpurun_out/lbvh_stress_$b.log 2>&1
  CRT_BVH_BUILDER=$b timeout 300 python tools/quick_perf.py rtiow 1920 1080 64 50 3 2 > gpurun_out/lbvh_rtiow_$b.log 2>&1
done
grep -E "SAH statistic|passed|failed|rror|refit" gpurun_out/pytest_lbvh.log | tail -30; for f in gpurun_out/lbvh_*.log; do echo $f; grep -h "crt\]\|scene\|Mrays" $f | tail -3; done
