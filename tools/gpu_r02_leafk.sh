#!/bin/bash
# experiment: the round's leaf visit only when >= K lanes wait at a leaf (libcrt_leafk.so, -DCRT_PIPE_LEAFK) against the committed build
mkdir -p gpurun_out; out=gpurun_out/${1:-r02k}_leafk.txt; : > $out
qp() { timeout 120 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays|sha1" | tail -2; }
A=/root/repo/cudaraytracer_b200/libcrt.so; B=/root/repo/cudaraytracer_b200/libcrt_leafk.so
for rep in 1 2; do echo "== head" >> $out; CRT_LIB_PATH=$A qp rtiow 1920 1080 64 50 3 3 >> $out
for k in 0 3 5 7 9; do echo "== leafk K=$k" >> $out; CRT_PIPE_LEAFK=$k CRT_LIB_PATH=$B qp rtiow 1920 1080 64 50 3 3 >> $out; done; done
for k in 5 7; do for t in 10 18; do echo "== leafk K=$k THRESH=$t" >> $out; CRT_PIPE_THRESH=$t CRT_PIPE_LEAFK=$k CRT_LIB_PATH=$B qp rtiow 1920 1080 64 50 3 2 >> $out; done; done
grep -v "^accum" $out; grep "^accum" $out | sort | uniq -c
