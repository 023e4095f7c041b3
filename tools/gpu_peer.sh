#!/bin/bash
# Peer-memory close-out visit: bash tools/gpu_peer.sh N TAG (run under gpurun --gpus N). Correctness first, then the timings.
N=${1:-2}
T=${2:-r01_m}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
t0=$SECONDS
timeout 120 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu_n$N.log 2>&1; echo "pytest rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_pytest_gpu_n$N.log
timeout 120 $TR --master-port 29521 tools/peer_check.py 1920 1080 8 rtiow > gpurun_out/${T}_peer_check_1080p_n$N.log 2>&1; echo "rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_peer_check_1080p_n$N.log
timeout 150 $TR --master-port 29522 tools/peer_check.py 7680 4320 $N stress > gpurun_out/${T}_peer_check_8k_n$N.log 2>&1; echo "rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_peer_check_8k_n$N.log
if [ "${3:-bench}" = "bench" ]; then
for c in peer nccl; do
timeout 150 $TR --master-port 29523 bench.py --gpus $N --steps 8 --warmup 3 --workload C5 --spp $((8 / N > 0 ? 8 / N : 1)) --no-cpu --no-refkernel --closeout $c > gpurun_out/${T}_bench_c5_n${N}_$c.log 2>&1; echo "rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_bench_c5_n${N}_$c.log
done
fi
tail -n 3 gpurun_out/${T}_pytest_gpu_n$N.log; grep -h "peer_check\|rc=" gpurun_out/${T}_peer_check_*_n$N.log | cut -c 1-300; for c in peer nccl; do tail -n 2 gpurun_out/${T}_bench_c5_n${N}_$c.log | cut -c 1-200; done
