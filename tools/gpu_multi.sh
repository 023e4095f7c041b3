#!/bin/bash
set -x
mkdir -p gpurun_out
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/bench_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 2 --warmup 3 --workload C4 --spp 32 --no-cpu --no-refkernel > gpurun_out/bench_c4_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_c4_n$N.log
timeout 600 python bench.py --gpus 1 --steps 2 --warmup 3 --workload C4 --spp 32 --no-cpu --no-refkernel > gpurun_out/bench_c4_n1.log 2>&1
timeout 600 python bench.py --impl reference --gpus $N --steps 1 --warmup 1 > gpurun_out/bench_ref_n$N.log 2>&1
