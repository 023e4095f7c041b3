#!/bin/bash
# which tree / variant combinations of the 1 M-sphere frame agree? (sha-1 of the accumulation buffer, 4K, 1 spp)
mkdir -p gpurun_out; out=gpurun_out/${1:-dbg}_sha.txt; : > $out
for b in lbvh host; do for v in 1 2 3; do
  echo "== builder $b variant $v" >> $out
  CRT_BVH_BUILDER=$b timeout 300 python tools/quick_perf.py grid 3840 2160 1 50 $v 1 2>&1 | grep -E "sha1|Mrays" | tail -2 >> $out
done; done
for h in 100 300; do for v in 1 3; do echo "== lbvh hint $h variant $v (CRT_GPU_BUILD_MIN_PRIMS=2)" >> $out; CRT_GPU_BUILD_MIN_PRIMS=2 CRT_BVH_BUILDER=lbvh timeout 300 python tools/quick_perf.py grid 3840 2160 1 50 $v 1 $h 2>&1 | grep -E "sha1" | tail -1 >> $out; done; done
cat $out
