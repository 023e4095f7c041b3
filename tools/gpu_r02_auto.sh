#!/bin/bash
# where is the megakernel / pipeline crossover now? (AUTO switches at 256 primitives, set in round 1)
mkdir -p gpurun_out; out=gpurun_out/${1:-r02a}_auto.txt; : > $out
qp() { timeout 120 python tools/quick_perf.py "$@" 2>&1 | grep -E "^scene|Mrays" | tail -2; }
for sc in grid stress; do for h in 6 8 10 12 14 16 20; do
  d=50; [ $sc = stress ] && d=100
  echo "== $sc hint $h mega" >> $out; qp $sc 1920 1080 32 $d 1 2 $h >> $out
  echo "== $sc hint $h pipe" >> $out; qp $sc 1920 1080 32 $d 3 2 $h | tail -1 >> $out
done; done
cat $out
