#!/bin/bash
# refill threshold / rounds per look of k_render_pipe after the ray-adoption and global-node steps: is (14, 16) / (1, 64) still the optimum?
T=${1:-r02th}; mkdir -p gpurun_out; out=gpurun_out/${T}_thresh.txt; : > $out
qp() { timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays" | tail -1; }
echo "== defaults rtiow 1080p 64 spp" >> $out; qp rtiow 1920 1080 64 50 3 2 >> $out
for t in 8 10 12 16 18 20; do echo "== THRESH=$t rtiow" >> $out; CRT_PIPE_THRESH=$t qp rtiow 1920 1080 64 50 3 2 >> $out; done
for r in 4 8 12 24 32; do echo "== ROUNDS=$r rtiow" >> $out; CRT_PIPE_ROUNDS=$r qp rtiow 1920 1080 64 50 3 2 >> $out; done
echo "== defaults grid 4K 4 spp" >> $out; qp grid 3840 2160 4 50 3 2 >> $out
echo "== defaults stress 4K 4 spp" >> $out; qp stress 3840 2160 4 100 3 2 >> $out
for t in 2 4 8; do
  echo "== THRESH=$t grid" >> $out; CRT_PIPE_THRESH=$t qp grid 3840 2160 4 50 3 2 >> $out
  echo "== THRESH=$t stress" >> $out; CRT_PIPE_THRESH=$t qp stress 3840 2160 4 100 3 2 >> $out
done
for r in 16 32 128; do
  echo "== ROUNDS=$r grid" >> $out; CRT_PIPE_ROUNDS=$r qp grid 3840 2160 4 50 3 2 >> $out
  echo "== ROUNDS=$r stress" >> $out; CRT_PIPE_ROUNDS=$r qp stress 3840 2160 4 100 3 2 >> $out
done
cat $out
