#!/bin/bash
# only the ncu launch list of the bench command (see gpu_final2.sh for why CRT_AUTO_CALIBRATE=0). $1 = tag.
T=${1:-r02_y4}; mkdir -p gpurun_out
CRT_AUTO_CALIBRATE=0 timeout 300 python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-refkernel --no-c4 > gpurun_out/${T}_bench_spp16_plain.log 2>&1 && \
CRT_AUTO_CALIBRATE=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_bench_spp16.csv python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-refkernel --no-c4 > gpurun_out/${T}_launches_ncu.log 2>&1
python tools/launch_summary.py gpurun_out/${T}_launches_bench_spp16.csv | head -14
grep '^{' gpurun_out/${T}_bench_spp16_plain.log | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['config']['variant_resolved'], d['gpu_launches'], d['ms_per_step'], d['roofline']['kernel_share_of_step'])"
