#!/bin/bash
# the megakernel on the PAIR staging (default) against the word-major table (CRT_MEGA_NO_PAIR=1), small scenes
mkdir -p gpurun_out; out=gpurun_out/${1:-r02m}_mega.txt; : > $out
qp() { timeout 120 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays|sha1" | tail -2; }
for rep in 1 2; do
for sc in "textured 1920 1080 64 50 1 2" "default 1920 1080 64 50 1 2" "rtiow 1920 1080 16 50 1 2"; do
  echo "== word-major: $sc" >> $out; CRT_MEGA_NO_PAIR=1 qp $sc >> $out
  echo "== PAIR: $sc" >> $out; qp $sc >> $out
done; done
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_reference_kernel.py -m gpu -x -q 2>&1 | tail -3 >> $out
cat $out
