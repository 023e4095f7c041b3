#!/bin/bash
# where between 486 and 1601 primitives does (20, 4) stop paying? grids of 401 / 785 / 1025 spheres, diffuse-heavy and glass / metal
T=${1:-r02th7}; mkdir -p gpurun_out; out=gpurun_out/${T}_thresh.txt; : > $out
qp() { timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "scene|Mrays" | tail -2; }
for h in 20 28 32 36; do for s in grid stress; do d=50; [ $s = stress ] && d=100
  for cfg in "14 16" "20 4" "18 5" "16 6"; do set -- $cfg
    echo "== $s hint $h THRESH=$1 ROUNDS=$2" >> $out; CRT_PIPE_THRESH=$1 CRT_PIPE_ROUNDS=$2 qp $s 1920 1080 16 $d 3 2 $h >> $out
  done
done; done
cat $out
