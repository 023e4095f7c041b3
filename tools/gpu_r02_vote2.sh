#!/bin/bash
# experiment: one vote per TWO node,node,node,leaf sub-rounds (libcrt_vote2.so, -DCRT_PIPE_VOTE2) against the committed build
mkdir -p gpurun_out; out=gpurun_out/${1:-r02v2}_vote2.txt; : > $out
qp() { timeout 120 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays|sha1" | tail -2; }
A=/root/repo/cudaraytracer_b200/libcrt.so; B=/root/repo/cudaraytracer_b200/libcrt_vote2.so
for rep in 1 2; do echo "== head" >> $out; CRT_LIB_PATH=$A qp rtiow 1920 1080 64 50 3 3 >> $out
echo "== vote2" >> $out; CRT_LIB_PATH=$B qp rtiow 1920 1080 64 50 3 3 >> $out; done
for r in 4 8 16; do for t in 12 14 18; do echo "== vote2 ROUNDS=$r THRESH=$t" >> $out; CRT_PIPE_ROUNDS=$r CRT_PIPE_THRESH=$t CRT_LIB_PATH=$B qp rtiow 1920 1080 64 50 3 2 >> $out; done; done
for lib in $A $B; do echo "== $(basename $lib) textured pipe" >> $out; CRT_LIB_PATH=$lib qp textured 1920 1080 64 50 3 2 >> $out; done
grep -v "^accum" $out; grep "^accum" $out | sort | uniq -c
