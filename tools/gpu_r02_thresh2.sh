#!/bin/bash
# second sweep: the refill threshold above 20 and the rounds per look below 8, alone and together (scenes staged whole in shared memory)
T=${1:-r02th2}; mkdir -p gpurun_out; out=gpurun_out/${T}_thresh.txt; : > $out
qp() { timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays" | tail -1; }
echo "== defaults rtiow 1080p 64 spp" >> $out; qp rtiow 1920 1080 64 50 3 2 >> $out
for t in 20 22 24 26 28 30 32; do for r in 16 4; do echo "== THRESH=$t ROUNDS=$r rtiow" >> $out; CRT_PIPE_THRESH=$t CRT_PIPE_ROUNDS=$r qp rtiow 1920 1080 64 50 3 2 >> $out; done; done
for r in 1 2 3 6; do echo "== THRESH=14 ROUNDS=$r rtiow" >> $out; CRT_PIPE_ROUNDS=$r qp rtiow 1920 1080 64 50 3 2 >> $out; done
for cfg in "14 16" "20 16" "24 4" "28 4"; do set -- $cfg
  echo "== THRESH=$1 ROUNDS=$2 rtiow 400x225 16 spp" >> $out; CRT_PIPE_THRESH=$1 CRT_PIPE_ROUNDS=$2 qp rtiow 400 225 16 50 3 3 >> $out
  echo "== THRESH=$1 ROUNDS=$2 textured 1080p 64 spp (pipeline)" >> $out; CRT_PIPE_THRESH=$1 CRT_PIPE_ROUNDS=$2 qp textured 1920 1080 64 50 3 2 >> $out
  echo "== THRESH=$1 ROUNDS=$2 rtiow 1080p 1 spp" >> $out; CRT_PIPE_THRESH=$1 CRT_PIPE_ROUNDS=$2 qp rtiow 1920 1080 1 50 3 3 >> $out
done
cat $out
