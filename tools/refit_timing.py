import time, numpy as np, sys
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cudaraytracer_b200 import Renderer, Scene, capi
s = Scene.generate(capi.SCENE_GRID_1M, 1984, 300)
h = s.hittable_array().copy()
r = Renderer(0)
r.set_bvh_builder(capi.BUILDER_HOST_SAH)
r.upload(s); r.set_camera(s.default_camera(160, 90)); r.resize(160, 90)
rng = np.random.default_rng(2)
for rep in range(4):
    for i in rng.choice(np.arange(1, len(h)), 2000, replace=False):
        i = int(i)
        r.edit_hittable(i, type=int(h["type"][i]), active=1, cx=float(h["cx"][i]), cy=float(h["cy"][i]) + float(rng.uniform(0, 3)), cz=float(h["cz"][i]),
                        a=float(h["a"][i]) * float(rng.uniform(0.8, 3.0)), b=0.0, material=int(h["material"][i]))
    t0 = time.perf_counter(); rb = r.commit(); t = time.perf_counter() - t0
    print("commit", rb, round(t * 1e3, 2), "ms; build_ms", r.bvh_stats().build_ms)
