#!/bin/bash
# end-of-launch rule of the pipelined kernel (partial batches once the work items are gone): small and large frames, A = libcrt_ab_head.so
mkdir -p gpurun_out; out=gpurun_out/${1:-r02t}_tail.txt; : > $out
qp() { timeout 120 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays|sha1" | tail -2; }
A=/root/repo/cudaraytracer_b200/libcrt_ab_head.so; B=/root/repo/cudaraytracer_b200/libcrt.so
for cfg in "rtiow 400 225 16 50" "rtiow 400 225 64 50" "rtiow 960 540 4 50" "rtiow 1920 1080 1 50" "rtiow 1920 1080 16 50" "rtiow 1920 1080 64 50" "stress 3840 2160 4 100" "grid 3840 2160 4 50" "stress 1280 720 1 100"; do
  for lib in $A $B; do echo "== $(basename $lib) $cfg" >> $out; CRT_LIB_PATH=$lib qp $cfg 3 5 >> $out; done
done
grep -v "^accum" $out; grep "^accum" $out | sort | uniq -c
