"""k_resolve_peers<G> on ONE GPU, for an ncu capture (profiling rules: one GPU, never a multi-rank command): G contexts on GPU 0
render their sample shards at W x H, then ONE call sums + resolves ALL rows over the G buffers (plain local pointers here, peer
mappings across real GPUs). The kernel's byte counts are what the capture is for: it must read 32 B per pixel per buffer and
write 4 B per pixel.   python tools/resolve_peers_one_gpu.py [G W H]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cudaraytracer_b200 import Renderer, Scene, capi
from cudaraytracer_b200.multigpu import sample_shard

G = int(sys.argv[1]) if len(sys.argv) > 1 else 8
W = int(sys.argv[2]) if len(sys.argv) > 2 else 3840
H = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
s = Scene.generate(capi.SCENE_RTIOW, 1984)
cam = s.default_camera(W, H)
ranks = []
for g in range(G):
    r = Renderer(0)
    r.upload(s); r.set_camera(cam); r.resize(W, H)
    b, st, c = sample_shard(g, G, G)
    r.render(spp=c, max_depth=8, spp_begin=b, spp_stride=st)
    r.synchronize()
    ranks.append(r)
ptrs = [r.accum_ptr() for r in ranks]
for _ in range(3):
    ranks[0].resolve_peers(ptrs, 0, H, ranks[0].rgba_ptr())
    ranks[0].synchronize()
img = ranks[0].read_rgba()
print(f"resolve_peers<{G}> {W}x{H}: frame complete {bool(((img >> 24) == 255).all())}; algorithmic bytes per launch: "
      f"{32 * G * W * H} read + {4 * W * H} written")
for r in ranks:
    r.close()
