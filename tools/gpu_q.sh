#!/bin/bash
mkdir -p gpurun_out
for v in 1 3; do timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 $v 2 > gpurun_out/q_rtiow_v$v.log 2>&1; done
CRT_BVH_BUILDER=host timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 3 2 > gpurun_out/q_grid_v3.log 2>&1
timeout 300 python tools/quick_perf.py stress 3840 2160 4 100 3 2 > gpurun_out/q_stress_v3.log 2>&1
timeout 300 python tools/quick_perf.py textured 1920 1080 16 50 1 2 > gpurun_out/q_textured_v1.log 2>&1
for f in gpurun_out/q_*.log; do echo $f; grep -h "scene\|Mrays" $f | tail -2; done
