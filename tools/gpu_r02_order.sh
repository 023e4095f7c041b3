#!/bin/bash
# Round-2 visit o: memory order of the nodes below the staged top of the tree (host builder): breadth-first throughout
# (CRT_PACK_TOP_BFS=1000000000, the previous layout) against depth-first subtrees below the first K breadth-first nodes.
T=${1:-r02o}; mkdir -p gpurun_out; out=gpurun_out/${T}_ab.txt; : > $out
qp() { timeout 300 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays|sha1" | tail -2; }
for k in 1000000000 4096 1024 512; do
  echo "== top_bfs $k stress 4K 4 spp depth 100" >> $out; CRT_PACK_TOP_BFS=$k qp stress 3840 2160 4 100 3 2 >> $out
  echo "== top_bfs $k stress 8K 2 spp depth 100" >> $out; CRT_PACK_TOP_BFS=$k qp stress 7680 4320 2 100 3 2 >> $out
  echo "== top_bfs $k grid 4K 4 spp host builder" >> $out; CRT_BVH_BUILDER=host CRT_PACK_TOP_BFS=$k qp grid 3840 2160 4 50 3 2 >> $out
done
echo "== grid 4K 4 spp lbvh (for comparison)" >> $out; qp grid 3840 2160 4 50 3 2 >> $out
cat $out
