#!/bin/bash
# 2-GPU visit with the final kernels: the whole gated GPU suite (incl. the tests that need >= 2 GPUs), then the driver's line at N = 2
T=${1:-r02_n2}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
grep -h "LaunchKernel drop-in" gpurun_out/${T}_pytest.log; tail -n 4 gpurun_out/${T}_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/${T}_bench_n2.json 2> gpurun_out/${T}_bench_n2.err
echo "bench rc=$?"; grep -v "OMP_NUM\|^\*\*\*\|^$" gpurun_out/${T}_bench_n2.err | tail -n 5; cut -c1-400 gpurun_out/${T}_bench_n2.json
