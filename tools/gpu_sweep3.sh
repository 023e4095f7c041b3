#!/bin/bash
# big-scene sweep of the pipeline's shared-memory split: pool size vs node prefix vs L1 (what is not shared memory)
mkdir -p gpurun_out
out=gpurun_out/sweep3.log; : > $out
for sc in "grid 3840 2160 4 50" "stress 3840 2160 4 100"; do
  for kb in 112 96 88 80 76 73; do
    for slots in 1024; do
      echo "== $sc CRT_SMEM_KB=$kb CRT_PIPE_SLOTS=$slots" >> $out
      CRT_SMEM_KB=$kb CRT_PIPE_SLOTS=$slots timeout 120 python tools/quick_perf.py $sc 3 2 2>&1 | grep -E "Mrays|rror" | tail -1 >> $out
    done
  done
done
cat $out
