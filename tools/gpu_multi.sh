#!/bin/bash
# Multi-GPU visit: bash tools/gpu_multi.sh N TAG  (run under gpurun --gpus N)
set -x
mkdir -p gpurun_out
N=${1:-2}
T=${2:-r01_f}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > gpurun_out/${T}_topo_n$N.txt 2>&1
timeout 900 $TR --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/${T}_bench_c2_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_c2_n$N.log
timeout 900 $TR --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --workload C4 --spp $((256 / N)) --no-cpu --no-refkernel > gpurun_out/${T}_bench_c4_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_c4_n$N.log
timeout 900 $TR --master-port 29513 bench.py --gpus $N --steps 8 --warmup 3 --workload C5 --spp $((8 / N > 0 ? 8 / N : 1)) --no-cpu --no-refkernel > gpurun_out/${T}_bench_c5_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_c5_n$N.log
timeout 600 $TR --master-port 29514 bench.py --impl reference --gpus $N --steps 1 --warmup 1 > gpurun_out/${T}_bench_ref_n$N.log 2>&1
timeout 600 python -m pytest tests/test_gpu_multigpu.py -m gpu -x -q > gpurun_out/${T}_pytest_multigpu_n$N.log 2>&1
tail -c 600 gpurun_out/${T}_bench_c2_n$N.log gpurun_out/${T}_bench_c4_n$N.log gpurun_out/${T}_bench_c5_n$N.log
