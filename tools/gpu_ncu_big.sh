#!/bin/bash
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe --launch-skip 1 -c 1 -o gpurun_out/r01_g_pipe_grid_full -f python tools/quick_perf.py grid 3840 2160 4 50 3 1 > gpurun_out/ncu_big.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe --launch-skip 1 -c 1 -o gpurun_out/r01_g_pipe_stress_full -f python tools/quick_perf.py stress 3840 2160 4 100 3 1 > gpurun_out/ncu_big2.log 2>&1
ls -la gpurun_out/*.ncu-rep
