#!/bin/bash
# third sweep: refill threshold x rounds per look around (14..22, 3..5)
T=${1:-r02th3}; mkdir -p gpurun_out; out=gpurun_out/${T}_thresh.txt; : > $out
qp() { timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays" | tail -1; }
for t in 12 14 16 18 20 22; do for r in 3 4 5; do echo "== THRESH=$t ROUNDS=$r rtiow" >> $out; CRT_PIPE_THRESH=$t CRT_PIPE_ROUNDS=$r qp rtiow 1920 1080 64 50 3 2 >> $out; done; done
for cfg in "14 16" "14 3" "18 3" "20 3" "20 4"; do set -- $cfg
  echo "== THRESH=$1 ROUNDS=$2 rtiow 400x225 16 spp" >> $out; CRT_PIPE_THRESH=$1 CRT_PIPE_ROUNDS=$2 qp rtiow 400 225 16 50 3 3 >> $out
  echo "== THRESH=$1 ROUNDS=$2 rtiow 1080p 1 spp" >> $out; CRT_PIPE_THRESH=$1 CRT_PIPE_ROUNDS=$2 qp rtiow 1920 1080 1 50 3 3 >> $out
  echo "== THRESH=$1 ROUNDS=$2 rtiow 1080p 16 spp" >> $out; CRT_PIPE_THRESH=$1 CRT_PIPE_ROUNDS=$2 qp rtiow 1920 1080 16 50 3 3 >> $out
  echo "== THRESH=$1 ROUNDS=$2 default scene 1080p 64 spp" >> $out; CRT_PIPE_THRESH=$1 CRT_PIPE_ROUNDS=$2 qp default 1920 1080 64 50 3 2 >> $out
done
cat $out
