#!/bin/bash
# Round 2 kernel iteration visit: bit-identity of the experimental schedules, then C2's scene at 64 spp for each env set given
# as arguments ("A=1 B=2" strings), default CRT_PIPE_SCHED=0 and 12.
out=gpurun_out/r02d.txt; : > $out
CRT_TEST_EXPERIMENTAL=1 timeout 300 python -m pytest tests -m gpu -q -k "experimental" >> $out 2>&1
run() { echo "== $*" >> $out; env "$@" timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 2 2>&1 | grep -E "Mrays|sha1" | tail -2 >> $out; }
if [ $# -eq 0 ]; then set -- "CRT_PIPE_SCHED=0" "CRT_PIPE_SCHED=12"; fi
for e in "$@"; do run $e; done
cat $out
