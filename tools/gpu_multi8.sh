#!/bin/bash
# 8-GPU visit (one call, 8x the box time): the driver's own line (C2 weak + strong-scaled C4), then C4 and C5 strong-scaled with
# each close-out, and the close-out timings of tools/peer_check.py at 8K. $1 = N, $2 = tag.
N=${1:-8}; T=${2:-r02p}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
run() { name=$1; shift; timeout 600 $TR bench.py --gpus $N "$@" > gpurun_out/${T}_${name}_n$N.json 2> gpurun_out/${T}_${name}_n$N.err; echo "$name rc=$?"; }
run default --steps 5 --warmup 3
for co in peer tiles; do run c4_$co --workload C4 --scaling strong --closeout $co --steps 5 --warmup 3 --no-cpu --no-refkernel; done
for co in nccl peer tiles; do run c5_$co --workload C5 --scaling strong --closeout $co --steps 5 --warmup 3 --no-cpu --no-refkernel; done
timeout 300 $TR tools/peer_check.py 7680 4320 8 stress > gpurun_out/${T}_peer_check_8k_n$N.txt 2>&1; echo "peer_check rc=$?"
grep -h "peer_check world" gpurun_out/${T}_peer_check_8k_n$N.txt
nvidia-smi topo -m > gpurun_out/${T}_topo_n$N.txt 2>&1
for f in gpurun_out/${T}_*_n$N.err; do echo $f; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" $f | tail -n 3; done
