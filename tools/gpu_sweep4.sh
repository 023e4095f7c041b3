#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 2 gpurun_out/pytest_gpu.log
out=gpurun_out/sweep4.log; : > $out
for sc in "grid 3840 2160 4 50" "stress 3840 2160 4 100"; do
  for pk in 8 24 64; do for sl in 1024 2048; do
    echo "== PREFIX_KB=$pk SLOTS=$sl $sc" >> $out
    CRT_PIPE_PREFIX_KB=$pk CRT_PIPE_SLOTS=$sl timeout 120 python tools/quick_perf.py $sc 3 2 2>&1 | grep Mrays | tail -1 >> $out
  done; done
done
echo "== default rtiow" >> $out; timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 2 2>&1 | grep Mrays | tail -1 >> $out
cat $out
