#!/bin/bash
# variant x scene matrix for DESIGN.md 3.1 (device-resident scene, quick_perf)
mkdir -p gpurun_out
out=gpurun_out/matrix.log; : > $out
for sc in "textured 1920 1080 16 50" "rtiow 1920 1080 64 50" "stress 3840 2160 4 100" "grid 3840 2160 4 50"; do
  for v in 1 2 3; do
    echo "== $sc variant $v" >> $out
    timeout 300 python tools/quick_perf.py $sc $v 2 2>&1 | grep -E "Mrays|rror" | tail -1 >> $out
  done
done
cat $out
