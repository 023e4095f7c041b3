#!/bin/bash
# bash tools/gpu_peer2.sh N TAG MODES [pytest]: C5 close-out comparison (nccl / peer / tiles) on N GPUs; optional GPU tests first.
N=${1:-2}
T=${2:-r01_n}
MODES=${3:-"peer tiles"}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
t0=$SECONDS
if [ "$4" = "pytest" ]; then
timeout 120 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu_n$N.log 2>&1; echo "pytest rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_pytest_gpu_n$N.log
tail -n 3 gpurun_out/${T}_pytest_gpu_n$N.log
fi
if [ "$5" != "nocheck" ]; then
timeout 150 $TR --master-port 29522 tools/peer_check.py 7680 4320 $N stress > gpurun_out/${T}_peer_check_8k_n$N.log 2>&1; echo "rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_peer_check_8k_n$N.log
grep -h "peer_check\|rc=\|Error\|error" gpurun_out/${T}_peer_check_8k_n$N.log | cut -c 1-400 | tail -5
fi
for c in $MODES; do
timeout 150 $TR --master-port 29523 bench.py --gpus $N --steps 8 --warmup 3 --workload C5 --spp $((8 / N > 0 ? 8 / N : 1)) --no-cpu --no-refkernel --closeout $c > gpurun_out/${T}_bench_c5_n${N}_$c.log 2>&1; echo "rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_bench_c5_n${N}_$c.log
tail -n 2 gpurun_out/${T}_bench_c5_n${N}_$c.log | cut -c 1-220
done
