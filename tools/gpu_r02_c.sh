#!/bin/bash
# Round 2, visit C: parameter sweep of the pipelined kernel around the node, node, node, leaf schedule (CRT_PIPE_SCHED=4) on RTIOW.
out=gpurun_out/r02c_sweep.txt; : > $out
run() { echo "== $*" >> $out; env "$@" timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 2 2>&1 | grep Mrays | tail -1 >> $out; }
run CRT_PIPE_SCHED=4
run CRT_PIPE_SCHED=4 CRT_PIPE_SLOTS=1024
run CRT_PIPE_SCHED=0 CRT_PIPE_SLOTS=1024
for t in 8 10 12 16 18 22; do run CRT_PIPE_SCHED=4 CRT_PIPE_THRESH=$t; done
for r in 4 8 32; do run CRT_PIPE_SCHED=4 CRT_PIPE_ROUNDS=$r; done
for c in 1 2 8 16; do run CRT_PIPE_SCHED=4 CRT_PIPE_CHECK=$c; done
for e in 0 64 1024; do run CRT_PIPE_SCHED=4 CRT_PIPE_EXT_LOW=$e; done
cat $out
