#!/bin/bash
# refill threshold x rounds per look on the staged scenes with deeper trees (1601 primitives), beside RTIOW
T=${1:-r02th5}; mkdir -p gpurun_out; out=gpurun_out/${T}_thresh.txt; : > $out
qp() { timeout 200 python tools/quick_perf.py "$@" 2>&1 | grep -E "Mrays" | tail -1; }
for t in 14 18 20; do for r in 4 6 8 16; do
  echo "== THRESH=$t ROUNDS=$r stress40" >> $out; CRT_PIPE_THRESH=$t CRT_PIPE_ROUNDS=$r qp stress 1920 1080 16 100 3 2 40 >> $out
  echo "== THRESH=$t ROUNDS=$r grid40" >> $out; CRT_PIPE_THRESH=$t CRT_PIPE_ROUNDS=$r qp grid 1920 1080 16 50 3 2 40 >> $out
  echo "== THRESH=$t ROUNDS=$r rtiow" >> $out; CRT_PIPE_THRESH=$t CRT_PIPE_ROUNDS=$r qp rtiow 1920 1080 64 50 3 2 >> $out
done; done
python - <<'P' >> $out
from cudaraytracer_b200 import capi, renderer
for name, kind, hint in (("rtiow", capi.SCENE_RTIOW, 0), ("stress40", capi.SCENE_STRESS_10K, 40), ("grid40", capi.SCENE_GRID_1M, 40), ("default", capi.SCENE_REFERENCE_DEFAULT, 0), ("textured", capi.SCENE_TEXTURED, 0)):
    r = renderer.Renderer(0)
    r.upload(renderer.Scene.generate(kind, size_hint=hint))
    st = r.bvh_stats()
    print("==", name, {f[0]: getattr(st, f[0]) for f in st._fields_})
P
cat $out
