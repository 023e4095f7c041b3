#!/bin/bash
# final one-GPU visit of round 2: gpu_final2.sh (tests, smoke, reference arm, benches, launch list, C2 capture) + captures of C3
T=${1:-r02_zz}
bash tools/gpu_final2.sh $T
timeout 300 python tools/quick_perf.py textured 1920 1080 64 50 1 1 > gpurun_out/${T}_c3_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render_mega -s 1 -c 1 -o gpurun_out/${T}_mega_c3_full -f python tools/quick_perf.py textured 1920 1080 64 50 1 1 > gpurun_out/${T}_mega_c3_ncu.log 2>&1
timeout 300 python tools/quick_perf.py grid 3840 2160 4 50 3 1 > gpurun_out/${T}_c4_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe -s 1 -c 1 -o gpurun_out/${T}_pipe_c4scene_full -f python tools/quick_perf.py grid 3840 2160 4 50 3 1 > gpurun_out/${T}_pipe_c4_ncu.log 2>&1
ls -la gpurun_out/${T}*.ncu-rep
