#!/bin/bash
# One full ncu capture of k_render_pipe on C2's scene at 128 spp. $1 = tag, remaining arguments = environment assignments.
T=$1; shift
env "$@" timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe -s 1 -c 1 -o gpurun_out/${T}_pipe_full -f python tools/quick_perf.py rtiow 1920 1080 128 50 3 1 > gpurun_out/${T}_ncu_full.log 2>&1
echo "rc=$?" >> gpurun_out/${T}_ncu_full.log; tail -5 gpurun_out/${T}_ncu_full.log; ls -la gpurun_out/${T}_pipe_full.ncu-rep
