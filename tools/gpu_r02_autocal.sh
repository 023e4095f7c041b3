#!/bin/bash
# measured AUTO (crt_last_variant): the parity file, then what AUTO picks and runs at on scenes either side of the crossover
mkdir -p gpurun_out; out=gpurun_out/${1:-r02ac}_autocal.txt; : > $out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3 >> $out
qp() { CRT_VERBOSE=1 timeout 120 python tools/quick_perf.py "$@" 2>&1 | grep -E "AUTO:|Mrays" | tail -3; }
for cfg in "grid 8" "grid 12" "stress 6" "stress 12" "textured 0" "default 0" "rtiow 0"; do
  set -- $cfg
  echo "== $1 hint $2: AUTO" >> $out; qp $1 1920 1080 64 50 0 2 $2 >> $out
  echo "== $1 hint $2: static rule (CRT_AUTO_CALIBRATE=0)" >> $out; CRT_AUTO_CALIBRATE=0 qp $1 1920 1080 64 50 0 2 $2 >> $out
done
timeout 300 python bench.py --workload C3 --steps 3 --warmup 3 --no-cpu --no-refkernel 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('bench C3', round(d['value']), d['config']['variant_resolved'], 'e2e', round(d['e2e']['value']))" >> $out
timeout 300 python bench.py --workload C1 --steps 20 --warmup 5 --no-cpu --no-refkernel 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('bench C1', round(d['value']), d['config']['variant_resolved'], 'e2e', round(d['e2e']['value']))" >> $out
cat $out
