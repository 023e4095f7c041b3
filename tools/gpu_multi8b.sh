#!/bin/bash
# lean 8-GPU visit with the final kernels: the driver's own line (C2 weak + strong-scaled C4), C5 with the tile close-out and with
# NCCL, the gated multi-GPU tests. $1 = N, $2 = tag.
N=${1:-8}; T=${2:-r02_zz8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
run() { name=$1; shift; timeout 600 $TR bench.py --gpus $N "$@" > gpurun_out/${T}_${name}_n$N.json 2> gpurun_out/${T}_${name}_n$N.err; echo "$name rc=$?"; }
run default --steps 5 --warmup 3
for co in tiles nccl; do run c5_$co --workload C5 --scaling strong --closeout $co --steps 5 --warmup 3 --no-cpu --no-refkernel; done
timeout 600 python -m pytest tests/test_gpu_multigpu.py -m gpu -x -q > gpurun_out/${T}_pytest_multigpu_n$N.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${T}_pytest_multigpu_n$N.log
for f in gpurun_out/${T}_*_n$N.err; do echo $f; grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$\|NCCL version" $f | tail -n 3; done
python - gpurun_out/${T}_default_n$N.json <<'P'
import sys, json
d = json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
c = d['c4_strong']
print('C2 weak', round(d['value']), 'e2e', round(d['e2e']['value']), '| C4 strong', round(c['value']), 'e2e', round(c['e2e']['value']), c['efficiency_vs_1gpu_same_run'], '|', d['parity'].get('multi_gpu_frame', {}).get('identical'))
P
