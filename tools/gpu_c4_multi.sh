#!/bin/bash
# C4 only: bash tools/gpu_c4_multi.sh N TAG
mkdir -p gpurun_out
N=${1:-8}
T=${2:-r01_i}
if [ "$N" = "1" ]; then
  timeout 900 python bench.py --workload C4 --steps 3 --warmup 3 > gpurun_out/${T}_bench_c4.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_c4.log
else
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --workload C4 --spp $((256 / N)) --no-cpu --no-refkernel > gpurun_out/${T}_bench_c4_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_c4_n$N.log
fi
