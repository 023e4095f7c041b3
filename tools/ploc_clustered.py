"""(Round 2: PLOC lost here too and was removed from the library; the script now compares the two builders that are left.)
Does PLOC ever beat the Morton LBVH? A CLUSTERED scene (what locally-ordered clustering is made for): Gaussian clusters of
spheres whose sizes span two decades, plus a ground sphere. Builds the tree with each builder, traces the same frame with the
pipelined kernel, prints build time and rate.   python tools/ploc_clustered.py [n_clusters per_cluster W H spp]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cudaraytracer_b200 import Renderer, Scene, capi

nc = int(sys.argv[1]) if len(sys.argv) > 1 else 300
per = int(sys.argv[2]) if len(sys.argv) > 2 else 700
W = int(sys.argv[3]) if len(sys.argv) > 3 else 1920
H = int(sys.argv[4]) if len(sys.argv) > 4 else 1080
spp = int(sys.argv[5]) if len(sys.argv) > 5 else 8
rng = np.random.default_rng(1984)
hit = [dict(type=capi.SPHERE, cx=0.0, cy=-1000.0, cz=0.0, a=1000.0, material=0)]
for c in range(nc):
    centre = np.array([rng.uniform(-60, 60), 0.0, rng.uniform(-60, 60)])
    spread = 10.0 ** rng.uniform(-0.5, 0.8)
    for k in range(per):
        r = float(10.0 ** rng.uniform(-2.2, -0.2) * spread * 0.2)
        p = centre + rng.normal(0, spread, 3) * np.array([1.0, 0.35, 1.0])
        hit.append(dict(type=capi.SPHERE, cx=float(p[0]), cy=float(abs(p[1]) + r), cz=float(p[2]), a=r, material=1 + (c + k) % 3))
mats = [dict(type=capi.LAMBERTIAN, texture=0), dict(type=capi.LAMBERTIAN, texture=1), dict(type=capi.METAL, texture=2, param=0.1),
        dict(type=capi.DIELECTRIC, texture=2, param=1.5)]
tex = [dict(type=capi.CONSTANT, color=(0.5, 0.5, 0.5)), dict(type=capi.CONSTANT, color=(0.7, 0.3, 0.2)),
       dict(type=capi.CONSTANT, color=(0.9, 0.9, 0.9))]
s = Scene.from_arrays(hit, mats, tex)
cam = Scene.generate(capi.SCENE_STRESS_10K, 1984, 4).default_camera(W, H)
cam.origin_x, cam.origin_y, cam.origin_z = 0.0, 45.0, 130.0
n = np.array([0.0, -0.33, -1.0]); n /= np.linalg.norm(n)
cam.orientation_x, cam.orientation_y, cam.orientation_z = (float(v) for v in n)
print(f"{len(hit)} spheres in {nc} clusters")
for name, b in (("host SAH", capi.BUILDER_HOST_SAH), ("GPU LBVH", capi.BUILDER_GPU_LBVH)) + ((("GPU PLOC", 3),) if os.environ.get("CRT_HAS_PLOC") else ()):
    r = Renderer(0)
    r.set_bvh_builder(b)
    r.upload(s)
    st = r.bvh_stats()
    r.set_camera(cam); r.resize(W, H)
    r.render(spp=1, max_depth=20, variant=capi.VARIANT_PIPELINE)
    best = 0.0
    for _ in range(3):
        r.clear(); r.reset_counters()
        r.render(spp=spp, max_depth=20, variant=capi.VARIANT_PIPELINE)
        ms = r.last_render_ms(); c = r.counters()
        best = max(best, c.rays / ms / 1e3)
    print(f"{name}: build {st.build_ms:.1f} ms, depth {st.max_depth}, SAH {st.sah_cost:.1f}, {best:.0f} Mrays/s, L {c.rays / c.paths:.2f}")
    r.close()
