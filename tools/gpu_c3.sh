#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload C3 --steps 2 --warmup 3 > gpurun_out/r01_e_bench_C3.log 2>&1; echo "rc=$?" >> gpurun_out/r01_e_bench_C3.log
tail -3 gpurun_out/pytest_gpu.log; tail -c 900 gpurun_out/r01_e_bench_C3.log
