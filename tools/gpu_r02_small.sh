#!/bin/bash
# small frames: where does the pipelined kernel's pool fill / drain cost more than the megakernel's divergence? (RTIOW, 486 spheres)
mkdir -p gpurun_out; out=gpurun_out/${1:-r02s}_small.txt; : > $out
qp() { timeout 120 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays" | tail -1; }
for cfg in "400 225 1" "400 225 4" "400 225 16" "400 225 64" "960 540 4" "960 540 16" "1920 1080 1" "1920 1080 4" "1920 1080 16"; do
  set -- $cfg
  echo "== mega $cfg" >> $out; qp rtiow $1 $2 $3 50 1 5 >> $out
  echo "== pipe $cfg" >> $out; qp rtiow $1 $2 $3 50 3 5 >> $out
done
cat $out
