#!/bin/bash
# GPU tests + device-side throughput of the pipelined kernel on the three scene classes (C2's, C4's and C5's scenes) and of
# AUTO on the small textured scene (C3's). Output: gpurun_out/<tag>_quick.txt. $1 = tag.
T=${1:-quick}
out=gpurun_out/${T}_quick.txt; : > $out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest_gpu.log
tail -n 3 gpurun_out/${T}_pytest_gpu.log >> $out
run() { echo "== $*" >> $out; timeout 300 python tools/quick_perf.py "$@" 2>&1 | grep -E "scene|Mrays|sha1" | tail -3 >> $out; }
run rtiow 1920 1080 64 50 3 2
run grid 3840 2160 4 50 3 2
run stress 3840 2160 4 100 3 2
run textured 1920 1080 64 50 0 2
run textured 1920 1080 64 50 3 2
run rtiow 400 225 16 50 0 3
cat $out
