#!/bin/bash
# Round-2 profiling visit (one GPU; every command has exited 0 without ncu before it runs under ncu). $1 = tag.
T=${1:-r02t}
mkdir -p gpurun_out
q() { timeout 300 python tools/quick_perf.py "$@"; }
# 1. the headline kernel: k_render_pipe on C2's scene, 128 spp per launch (-s 1 skips the 1-spp warm-up launch)
q rtiow 1920 1080 128 50 3 1 > gpurun_out/${T}_c2_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe -s 1 -c 1 -o gpurun_out/${T}_pipe_c2_full -f python tools/quick_perf.py rtiow 1920 1080 128 50 3 1 > gpurun_out/${T}_pipe_c2_ncu.log 2>&1
# 2. C3: the textured scene with its 8192x4096 image, AUTO (megakernel) and the pipelined kernel, 64 spp per launch
q textured 1920 1080 64 50 0 1 > gpurun_out/${T}_c3_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render_mega -s 1 -c 1 -o gpurun_out/${T}_mega_c3_full -f python tools/quick_perf.py textured 1920 1080 64 50 0 1 > gpurun_out/${T}_mega_c3_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render_pipe -s 1 -c 1 -o gpurun_out/${T}_pipe_c3_full -f python tools/quick_perf.py textured 1920 1080 64 50 3 1 > gpurun_out/${T}_pipe_c3_ncu.log 2>&1
# 3. the peer close-out kernel with 8 buffers at 4K on one GPU
timeout 300 python tools/resolve_peers_one_gpu.py 8 3840 2160 > gpurun_out/${T}_resolve_peers_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none -k regex:k_resolve_peers -s 2 -c 1 -o gpurun_out/${T}_resolve_peers8_full -f python tools/resolve_peers_one_gpu.py 8 3840 2160 > gpurun_out/${T}_resolve_peers_ncu.log 2>&1
# 4. launch list of the bench command (cold-cache, serialised: only the kernel's SHARE of the step is comparable)
timeout 300 python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-refkernel --no-c4 > gpurun_out/${T}_bench_spp16_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_bench_spp16.csv python bench.py --steps 1 --warmup 3 --spp 16 --no-cpu --no-refkernel --no-c4 > gpurun_out/${T}_launches_ncu.log 2>&1
cat gpurun_out/${T}_resolve_peers_plain.log; grep -h Mrays gpurun_out/${T}_c2_plain.log gpurun_out/${T}_c3_plain.log | tail -2; ls -la gpurun_out/*.ncu-rep
