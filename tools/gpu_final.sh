#!/bin/bash
# Round close-out visit: GPU tests, smoke, every BASELINE.json config on one GPU, the reference arm, ncu launch list and
# one full capture of the dominant kernel on the headline workload. Tag = $1 (file-name suffix).
set -x
T=${1:-r01_e}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
nproc > gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/${T}_bench_c2.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_c2.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.log 2>&1
for w in C1 C3 C4 C5; do timeout 900 python bench.py --workload $w --steps 2 --warmup 3 > gpurun_out/${T}_bench_$w.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_$w.log; done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-refkernel > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe -c 1 -o gpurun_out/${T}_pipe_full -f python bench.py --steps 1 --warmup 3 --spp 128 --no-cpu --no-refkernel > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -30
