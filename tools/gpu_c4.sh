#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload C4 --steps 2 --warmup 3 --no-cpu --no-refkernel > gpurun_out/c4_auto.log 2>&1
CRT_BVH_BUILDER=host timeout 600 python bench.py --workload C4 --steps 2 --warmup 3 --no-cpu --no-refkernel > gpurun_out/c4_host.log 2>&1
tail -n 3 gpurun_out/pytest_gpu.log
