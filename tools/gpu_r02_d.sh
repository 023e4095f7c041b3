#!/bin/bash
# Round 2, visit D: the paired-layout FFMA2 node visit (CRT_PIPE_SCHED bit 3) against the earlier schedules; bit-identity first.
out=gpurun_out/r02d.txt; : > $out
CRT_TEST_EXPERIMENTAL=1 timeout 300 python -m pytest tests -m gpu -q -k "experimental" >> $out 2>&1
run() { echo "== $*" >> $out; env "$@" timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 2 2>&1 | grep -E "Mrays|sha1" | tail -2 >> $out; }
for s in 0 4 8 9 12; do run CRT_PIPE_SCHED=$s; done
cat $out
