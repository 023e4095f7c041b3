#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
for v in 1 3 2; do timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 $v 2 > gpurun_out/quick_grid_v$v.log 2>&1; done
for v in 1 3; do timeout 300 python tools/quick_perf.py stress 3840 2160 4 100 $v 2 > gpurun_out/quick_stress4k_v$v.log 2>&1; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-refkernel > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe -c 1 -o gpurun_out/pipe_c2_full -f python bench.py --steps 1 --warmup 3 --spp 128 --no-cpu --no-refkernel > gpurun_out/ncu_full_c2.log 2>&1
ls -la gpurun_out
