#!/bin/bash
# Round 2, visit A: the experiments round 1 left queued (tools/gpu_quick3.sh) + compute-sanitizer on the three variants.
bash tools/gpu_quick3.sh > gpurun_out/r02a_quick3.log 2>&1
for v in 1 2 3; do
  for tool in memcheck racecheck; do
    timeout 400 compute-sanitizer --tool $tool --print-limit 20 python tools/quick_perf.py rtiow 96 54 2 50 $v 1 > gpurun_out/r02a_san_${tool}_v$v.log 2>&1
    echo "rc=$?" >> gpurun_out/r02a_san_${tool}_v$v.log
  done
done
cat gpurun_out/r02a_quick3.log | tail -40
for f in gpurun_out/r02a_san_*.log; do echo $f; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|rc=" $f; done
