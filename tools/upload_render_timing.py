"""Render time right after each upload (does a re-uploaded scene render as fast as the first one?)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cudaraytracer_b200 import Renderer, Scene, capi
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
s = Scene.generate(capi.SCENE_GRID_1M, 1984, 0)
r = Renderer(0)
W, H = 3840, 2160
cam = s.default_camera(W, H)
r.upload(s); r.set_camera(cam); r.resize(W, H)
for j in range(3):
    r.clear(); r.render(spp=spp, max_depth=50)
    print("resident render", j, round(r.last_render_ms(), 2), "ms", flush=True)
for k in range(3):
    t0 = time.perf_counter(); r.upload(s); t1 = time.perf_counter()
    r.set_camera(cam); r.clear(); r.render(spp=spp, max_depth=50)
    ms = r.last_render_ms(); t2 = time.perf_counter()
    print("upload", round((t1 - t0) * 1e3, 1), "ms; render", round(ms, 2), "ms; wall of both", round((t2 - t0) * 1e3, 1), "ms", flush=True)
