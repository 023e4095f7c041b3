#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 3 > gpurun_out/q3_rtiow.log 2>&1
timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 3 2 > gpurun_out/q3_grid.log 2>&1
timeout 300 python tools/quick_perf.py stress 3840 2160 4 100 3 2 > gpurun_out/q3_stress.log 2>&1
for sc in "rtiow 1920 1080 64 50" "grid 3840 2160 4 50" "stress 3840 2160 4 100"; do CRT_PIPE_SCHED=1 timeout 300 python tools/quick_perf.py $sc 3 2 > gpurun_out/q3_sched1_${sc%% *}.log 2>&1; done
for sc in "grid 3840 2160 4 50" "stress 3840 2160 4 100"; do CRT_PIPE_SCHED=4 timeout 300 python tools/quick_perf.py $sc 3 2 > gpurun_out/q3_sched4_${sc%% *}.log 2>&1; done
for e in 2 3 4 6; do CRT_PIPE_SCHED=$e timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 3 > gpurun_out/q3_sched${e}_rtiow.log 2>&1; done
CRT_TEST_EXPERIMENTAL=1 timeout 300 python -m pytest tests -m gpu -q -k "experimental or ipc_export or native_multigpu" > gpurun_out/q3_pytest_experimental.log 2>&1
for sc in "rtiow 1920 1080 64 50" "textured 1920 1080 64 50"; do CRT_BVH_ROTATE=4 timeout 120 python tools/quick_perf.py $sc 0 3 > gpurun_out/q3_rotate_${sc%% *}.log 2>&1; timeout 120 python tools/quick_perf.py $sc 0 3 > gpurun_out/q3_norotate_${sc%% *}.log 2>&1; done
CRT_PIPE_SLOTS=512 timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 2 > gpurun_out/q3_rtiow_generic.log 2>&1
tail -n 2 gpurun_out/pytest_gpu.log; for f in gpurun_out/q3_*.log; do echo $f; grep -h "Mrays" $f | tail -1; done
