#!/bin/bash
# N-GPU visit: the multi-GPU tests of the gated suite, then bench.py under torchrun (C2 weak as `value` + the strong-scaled C4 leg).
# $1 = N, $2 = tag, $3.. = extra bench.py arguments.
N=$1; T=$2; shift 2
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multigpu.py -m gpu -x -q > gpurun_out/${T}_pytest_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest_n$N.log
tail -n 3 gpurun_out/${T}_pytest_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 "$@" > gpurun_out/${T}_bench_n$N.json 2> gpurun_out/${T}_bench_n$N.err
echo "bench rc=$?"; tail -n 5 gpurun_out/${T}_bench_n$N.err; cut -c1-600 gpurun_out/${T}_bench_n$N.json
