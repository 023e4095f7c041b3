#!/bin/bash
# One gpurun visit: GPU tests, bench, ncu launch list, one full ncu capture of the dominant kernel.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
nproc > gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python tools/quick_perf.py rtiow 1920 1080 64 50 > gpurun_out/quick_rtiow.log 2>&1
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-refkernel > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_mega -c 1 -o gpurun_out/mega_full -f python tools/quick_perf.py rtiow 1920 1080 8 50 0 1 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
