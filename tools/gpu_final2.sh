#!/bin/bash
# Round-2 close-out visit on one GPU: GPU tests, smoke, the default bench (C2 + strong C4 leg), the reference arm, the other
# configs, the ncu launch list of the bench command and one full capture of k_render_pipe on C2's scene. $1 = tag.
T=${1:-r02_z}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${T}_smi.txt; nproc >> gpurun_out/${T}_smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${T}_smoke.log
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.log 2>&1
timeout 900 python bench.py > gpurun_out/${T}_bench_default.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_default.log
for w in C1 C3 C5; do timeout 900 python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/${T}_bench_$w.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_bench_$w.log; done
# The launch list: AUTO's measured choice (crt_render times both kernels on the first samples) is switched off for BOTH runs of
# this command: under ncu every launch is serialised and carries the tool's overhead, the timing flips to the megakernel and
# the list would show a kernel the timed step does not run. With the static rule the step is the pipelined kernel, as in bench.py.
CRT_AUTO_CALIBRATE=0 timeout 300 python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-refkernel --no-c4 > gpurun_out/${T}_bench_spp16_plain.log 2>&1 && \
CRT_AUTO_CALIBRATE=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_bench_spp16.csv python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-refkernel --no-c4 > gpurun_out/${T}_launches_ncu.log 2>&1
timeout 300 python tools/quick_perf.py rtiow 1920 1080 128 50 3 1 > gpurun_out/${T}_c2_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe -s 1 -c 1 -o gpurun_out/${T}_pipe_c2_full -f python tools/quick_perf.py rtiow 1920 1080 128 50 3 1 > gpurun_out/${T}_pipe_c2_ncu.log 2>&1
tail -5 gpurun_out/${T}_pytest_gpu.log; tail -2 gpurun_out/${T}_smoke.log
for f in default C1 C3 C5; do python - gpurun_out/${T}_bench_$f.log <<'P'
import sys, json
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
    print(sys.argv[1], round(d['value'], 1), d['unit'], 'ms', round(d['ms_per_step'], 3), 'frac', round(d['roofline']['frac'], 4), 'e2e', round(d['e2e']['value'], 1),
          'refk', (d.get('reference_kernel') or {}).get('ratio_new_over_reference_kernel'), 'c4', (d.get('c4_strong') or {}).get('value'))
except Exception as e:
    print(sys.argv[1], 'no line:', e)
P
done
ls -la gpurun_out/*.ncu-rep
