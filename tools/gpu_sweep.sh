#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
CRT_BVH_TIMING=1 timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 3 2 > gpurun_out/quick_grid_default.log 2>&1
for sl in 512 1024 2048; do CRT_PIPE_SLOTS=$sl timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 3 2 > gpurun_out/quick_grid_s$sl.log 2>&1; done
for sl in 512 1024 2048; do CRT_PIPE_SLOTS=$sl timeout 300 python tools/quick_perf.py stress 3840 2160 4 100 3 2 > gpurun_out/quick_stress_s$sl.log 2>&1; done
for t in 12 16 24; do CRT_PIPE_THRESH=$t timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 3 2 > gpurun_out/quick_grid_t$t.log 2>&1; done
timeout 600 python bench.py --gpus 1 --steps 2 --warmup 3 --workload C4 --spp 32 --no-cpu --no-refkernel > gpurun_out/bench_c4_n1.log 2>&1
timeout 600 python bench.py --gpus 1 --steps 2 --warmup 3 --workload C3 --no-cpu > gpurun_out/bench_c3_n1.log 2>&1
timeout 600 python bench.py --gpus 1 --steps 2 --warmup 3 --workload C5 --no-cpu --no-refkernel > gpurun_out/bench_c5_n1.log 2>&1
timeout 600 python bench.py --gpus 1 --steps 2 --warmup 3 --workload C1 > gpurun_out/bench_c1_n1.log 2>&1
