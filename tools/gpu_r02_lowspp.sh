#!/bin/bash
# few paths per launch: where does the megakernel beat the pipelined kernel (whose pool of 148 x 2048 paths never refills then)?
T=${1:-r02ls}; mkdir -p gpurun_out; out=gpurun_out/${T}_lowspp.txt; : > $out
qp() { timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays" | tail -1; }
for fr in "400 225" "960 540" "1920 1080"; do for spp in 1 2 4 8; do for sc in rtiow "stress H 20" "grid H 28"; do
  set -- $sc; h=""; d=50; [ "$2" = H ] && h=$3; [ $1 = stress ] && d=100
  for v in 1 3; do echo "== $1 $h ${fr/ /x} spp $spp variant $v" >> $out; qp $1 $fr $spp $d $v 5 $h >> $out; done
done; done; done
cat $out
