#!/bin/bash
mkdir -p gpurun_out
for ck in 1 4 16; do for lo in 64 256 1024; do for t in 20 24; do
  echo "check $ck ext_low $lo thresh $t: $(CRT_PIPE_CHECK=$ck CRT_PIPE_EXT_LOW=$lo CRT_PIPE_THRESH=$t timeout 120 python tools/quick_perf.py rtiow 1920 1080 64 50 3 2 2>&1 | grep Mrays | tail -1)"
done; done; done > gpurun_out/sweep2.log 2>&1
