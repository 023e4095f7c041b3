#!/bin/bash
# does the staged prefix of the node table still pay on the scenes that do not fit in shared memory?
mkdir -p gpurun_out; out=gpurun_out/${1:-r02pf}_prefix.txt; : > $out
qp() { timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays" | tail -1; }
for kb in ${KBS:-0 8 24 64}; do
  echo "== CRT_PIPE_PREFIX_KB=$kb grid 4K 4 spp" >> $out; CRT_PIPE_PREFIX_KB=$kb qp grid 3840 2160 4 50 3 2 >> $out
  echo "== CRT_PIPE_PREFIX_KB=$kb stress 4K 4 spp depth 100" >> $out; CRT_PIPE_PREFIX_KB=$kb qp stress 3840 2160 4 100 3 2 >> $out
done
cat $out
