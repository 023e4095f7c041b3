#!/bin/bash
# Short close-out visit for a nearly spent GPU budget: each leg has its own timeout and writes its own log, most valuable
# first (headline bench line, GPU tests, ncu launch list, one full capture of the dominant kernel). Tag = $1.
T=${1:-r01_l}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${T}_smi.txt
nproc > gpurun_out/${T}_nproc.txt
t0=$SECONDS
timeout 240 python bench.py --steps 3 --warmup 3 > gpurun_out/${T}_bench_c2.log 2> gpurun_out/${T}_bench_c2.err; echo "rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_bench_c2.err
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_pytest_gpu.log
timeout 150 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-refkernel > gpurun_out/${T}_ncu_launch.log 2>&1; echo "launches rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_ncu_launch.log
timeout 240 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe -c 1 -o gpurun_out/${T}_pipe_full -f python bench.py --steps 1 --warmup 3 --spp 128 --no-cpu --no-refkernel > gpurun_out/${T}_ncu_full.log 2>&1; echo "full rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_ncu_full.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_smoke.log
timeout 120 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.log 2>&1; echo "ref rc=$? t=$((SECONDS-t0))" >> gpurun_out/${T}_bench_ref.log
tail -n 3 gpurun_out/${T}_pytest_gpu.log; cat gpurun_out/${T}_bench_c2.log | cut -c 1-400; ls -la gpurun_out | tail -20
