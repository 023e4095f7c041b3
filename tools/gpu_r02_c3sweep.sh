#!/bin/bash
# C3's scene (textured, 67 primitives) on the pipelined kernel: refill threshold / rounds per refill / pool size, against AUTO (megakernel)
mkdir -p gpurun_out; out=gpurun_out/${1:-r02c3}_sweep.txt; : > $out
qp() { timeout 120 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays" | tail -1; }
echo "== megakernel (AUTO)" >> $out; qp textured 1920 1080 64 50 0 2 >> $out
echo "== pipeline default" >> $out; qp textured 1920 1080 64 50 3 2 >> $out
for t in 8 18 22 26 30; do echo "== pipeline THRESH=$t" >> $out; CRT_PIPE_THRESH=$t qp textured 1920 1080 64 50 3 2 >> $out; done
for r in 1 2 4 8; do echo "== pipeline ROUNDS=$r" >> $out; CRT_PIPE_ROUNDS=$r qp textured 1920 1080 64 50 3 2 >> $out; done
for r in 1 2; do for t in 22 28; do echo "== pipeline ROUNDS=$r THRESH=$t" >> $out; CRT_PIPE_ROUNDS=$r CRT_PIPE_THRESH=$t qp textured 1920 1080 64 50 3 2 >> $out; done; done
echo "== pipeline SLOTS=1024" >> $out; CRT_PIPE_SLOTS=1024 qp textured 1920 1080 64 50 3 2 >> $out
echo "== pipeline SLOTS=4096 (fits? 67 prims)" >> $out; CRT_PIPE_SLOTS=4096 qp textured 1920 1080 64 50 3 2 >> $out
echo "== default scene mega / pipe" >> $out; qp default 1920 1080 64 50 0 2 >> $out; qp default 1920 1080 64 50 3 2 >> $out
cat $out
