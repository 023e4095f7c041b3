"""Quick device-side throughput probe (not the bench): python tools/quick_perf.py [scene] [W H spp depth variant reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cudaraytracer_b200 import Renderer, Scene, capi

kinds = {"rtiow": (capi.SCENE_RTIOW, 0), "textured": (capi.SCENE_TEXTURED, 0), "grid": (capi.SCENE_GRID_1M, 0),
         "stress": (capi.SCENE_STRESS_10K, 0), "default": (capi.SCENE_REFERENCE_DEFAULT, 0)}
name = sys.argv[1] if len(sys.argv) > 1 else "rtiow"
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
H = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
spp = int(sys.argv[4]) if len(sys.argv) > 4 else 64
depth = int(sys.argv[5]) if len(sys.argv) > 5 else 50
variant = int(sys.argv[6]) if len(sys.argv) > 6 else 0
reps = int(sys.argv[7]) if len(sys.argv) > 7 else 3
hint = int(sys.argv[8]) if len(sys.argv) > 8 else kinds[name][1]
flags = int(sys.argv[9]) if len(sys.argv) > 9 else 0

s = Scene.generate(kinds[name][0], 1984, hint)
r = Renderer(0)
r.upload(s)
st = r.bvh_stats()
print(f"scene {name}: {st.n_prims} prims, {st.n_nodes} nodes, depth {st.max_depth}, SAH {st.sah_cost:.1f}, build {st.build_ms:.1f} ms")
r.set_camera(s.default_camera(W, H))
r.resize(W, H)
r.render(spp=1, max_depth=depth, variant=variant, flags=flags)
r.synchronize()
for i in range(reps):
    r.clear()
    r.reset_counters()
    r.render(spp=spp, max_depth=depth, variant=variant, flags=flags)
    ms = r.last_render_ms()
    c = r.counters()
    extra = f" box/ray {c.box_tests / c.rays:.1f} prim/ray {c.prim_tests / c.rays:.2f}" if flags & 2 else ""
    print(f"{W}x{H} spp {spp} depth {depth} variant {variant}: {ms:.2f} ms  {c.rays / ms / 1e3:.1f} Mrays/s  "
          f"{spp / ms * 1e3 * (W * H) / (1920 * 1080):.1f} 1080p-spp/s  L {c.rays / c.paths:.2f}{extra}")

import hashlib
acc = r.read_accum()
print("accum sha1", hashlib.sha1(acc.tobytes()).hexdigest()[:16], "mean", float(acc[..., :3].mean() / spp))
