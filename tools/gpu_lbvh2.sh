#!/bin/bash
mkdir -p gpurun_out
CRT_BVH_TIMING=1 CRT_BVH_BUILDER=gpu timeout 300 python - > gpurun_out/lbvh_timing.log 2>&1 <<'PY'
import time
from cudaraytracer_b200 import Renderer, Scene, capi
s = Scene.generate(capi.SCENE_GRID_1M, 1984, 0)
r = Renderer(0)
for i in range(4):
    t = time.time(); r.upload(s); print("upload", round((time.time() - t) * 1e3, 1), "ms", r.bvh_stats().build_ms)
PY
cat gpurun_out/lbvh_timing.log
