#!/bin/bash
# A/B of two builds of libcrt.so in ONE call (boxes differ between calls): $1 = path of build A, $2 = path of build B, alternating.
A=$1; B=$2; out=gpurun_out/ab.txt; : > $out
for rep in 1 2; do for lib in $A $B; do
  echo "== $lib rtiow 1080p 64 spp" >> $out; CRT_LIB_PATH=$lib timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 3 2>&1 | grep Mrays | tail -2 >> $out
done; done
for lib in $A $B; do
  echo "== $lib bench C1" >> $out; CRT_LIB_PATH=$lib timeout 200 python bench.py --workload C1 --steps 20 --warmup 5 --no-cpu --no-refkernel 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('C1 ms/step', round(d['ms_per_step'],4), 'value', round(d['value']), 'e2e ms', round(d['e2e']['ms_per_step'],4), 'launches', d['gpu_launches'])" >> $out
  echo "== $lib grid 4K" >> $out; CRT_LIB_PATH=$lib timeout 200 python tools/quick_perf.py grid 3840 2160 4 50 3 2 2>&1 | grep Mrays | tail -1 >> $out
done
cat $out
