#!/bin/bash
# Round-2 visit h3: binary16 half-extents in the GLOBAL node table of k_render_pipe, with EVERY fast path culling on the same
# (binary16-rounded) boxes. (1) do all tree x variant combinations of the 1 M-sphere frame agree, (2) the whole GPU suite,
# (3) A/B against the committed build (libcrt_ab_head.so) in the same call.
T=${1:-r02h3}
mkdir -p gpurun_out
bash tools/gpu_r02_dbg.sh ${T} > /dev/null 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/${T}_pytest.log
out=gpurun_out/${T}_ab.txt; : > $out
tail -4 gpurun_out/${T}_pytest.log >> $out
A=/root/repo/cudaraytracer_b200/libcrt_ab_head.so; B=/root/repo/cudaraytracer_b200/libcrt.so
qp() { timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays|sha1" | tail -2; }
for rep in 1 2; do for lib in $A $B; do
  echo "== $(basename $lib) rtiow 1080p 64 spp pipeline" >> $out; CRT_LIB_PATH=$lib qp rtiow 1920 1080 64 50 3 2 >> $out
  echo "== $(basename $lib) grid 4K 4 spp" >> $out; CRT_LIB_PATH=$lib qp grid 3840 2160 4 50 3 2 >> $out
  echo "== $(basename $lib) stress 4K 4 spp depth 100" >> $out; CRT_LIB_PATH=$lib qp stress 3840 2160 4 100 3 2 >> $out
done; done
for lib in $A $B; do
  echo "== $(basename $lib) stress 8K 2 spp depth 100" >> $out; CRT_LIB_PATH=$lib qp stress 7680 4320 2 100 3 2 >> $out
  echo "== $(basename $lib) rtiow 1080p 64 spp mega" >> $out; CRT_LIB_PATH=$lib qp rtiow 1920 1080 64 50 1 2 >> $out
  echo "== $(basename $lib) stress 4K 4 spp mega" >> $out; CRT_LIB_PATH=$lib qp stress 3840 2160 4 100 1 1 >> $out
done
cat gpurun_out/${T}_sha.txt; cat $out
