#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for v in 1 3; do timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 $v > gpurun_out/quick_rtiow_v$v.log 2>&1; done
for t in 12 16 24; do CRT_PIPE_THRESH=$t timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 > gpurun_out/quick_pipe_t$t.log 2>&1; done
for ck in 1 2; do CRT_PIPE_CHECK=$ck timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 > gpurun_out/quick_pipe_c$ck.log 2>&1; done
for sc in textured stress; do for v in 1 3; do timeout 200 python tools/quick_perf.py $sc 1920 1080 16 50 $v > gpurun_out/quick_${sc}_v$v.log 2>&1; done; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe --launch-skip 1 -c 1 -o gpurun_out/pipe_full -f python tools/quick_perf.py rtiow 1920 1080 16 50 3 1 > gpurun_out/ncu_full_pipe.log 2>&1
ls -la gpurun_out
