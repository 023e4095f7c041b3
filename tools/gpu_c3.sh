#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -X faulthandler bench.py --workload C3 --steps 1 --warmup 3 --no-cpu --no-refkernel > gpurun_out/c3_a.log 2>&1; echo "rc=$?" >> gpurun_out/c3_a.log
timeout 600 python -X faulthandler bench.py --workload C3 --steps 1 --warmup 3 --no-refkernel > gpurun_out/c3_b.log 2>&1; echo "rc=$?" >> gpurun_out/c3_b.log
timeout 600 python -X faulthandler bench.py --workload C3 --steps 1 --warmup 3 --no-cpu > gpurun_out/c3_c.log 2>&1; echo "rc=$?" >> gpurun_out/c3_c.log
timeout 600 python bench.py --workload C4 --steps 2 --warmup 3 > gpurun_out/r01_e_bench_C4.log 2>&1; echo "rc=$?" >> gpurun_out/r01_e_bench_C4.log
tail -c 1500 gpurun_out/c3_a.log gpurun_out/c3_b.log gpurun_out/c3_c.log
