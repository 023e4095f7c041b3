#!/bin/bash
# AUTO now also times the refill pair (20, 4) of the pipelined kernel beside the static (16, 6): what does it choose per scene, and
# what does the static pair give when the variant is named? $1 = tag.
T=${1:-r02ap}; mkdir -p gpurun_out; out=gpurun_out/${T}_autopair.txt; : > $out
qp() { CRT_VERBOSE=1 timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "AUTO|Mrays|sha1" | tail -4; }
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/${T}_pytest_parity.log 2>&1; tail -2 gpurun_out/${T}_pytest_parity.log >> $out
for s in "rtiow 1920 1080 64 50" "rtiow 1920 1080 16 50" "stress 1920 1080 16 100 H 20" "stress 1920 1080 16 100 H 28" "grid 1920 1080 16 50 H 28" "grid 1920 1080 16 50 H 36" "stress 1920 1080 16 100 H 40" "textured 1920 1080 64 50" "default 1920 1080 64 50"; do
  set -- $s; h=""; [ "$6" = H ] && h=$7
  echo "== AUTO: $s" >> $out; qp $1 $2 $3 $4 $5 0 3 $h >> $out
  echo "== pipeline, static pair: $s" >> $out; qp $1 $2 $3 $4 $5 3 2 $h >> $out
done
cat $out
