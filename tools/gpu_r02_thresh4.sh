#!/bin/bash
# the new defaults (THRESH 20, ROUNDS 4) against the old ones (14, 16) on every scene class the pipelined kernel stages whole
T=${1:-r02th4}; mkdir -p gpurun_out; out=gpurun_out/${T}_ab.txt; : > $out
qp() { timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays|sha1" | tail -2; }
for s in "rtiow 1920 1080 64 50 3 2" "rtiow 400 225 16 50 3 3" "rtiow 1920 1080 4 50 3 3" "stress 1920 1080 16 100 3 2 40" "grid 1920 1080 16 50 3 2 40" "default 1920 1080 64 50 3 2" "textured 1920 1080 64 50 3 2"; do
  echo "== old (14,16): $s" >> $out; CRT_PIPE_THRESH=14 CRT_PIPE_ROUNDS=16 qp $s >> $out
  echo "== new (20,4):  $s" >> $out; qp $s >> $out
done
cat $out
