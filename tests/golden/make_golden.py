"""Generates tests/golden/*.npz from the REFERENCE'S OWN headers compiled for the host
(oracle/_ref/libcrt_refhost.so; needs /root/reference to build it). Run in the authoring container:
    python tests/golden/make_golden.py
Each fixture: 64x36, 2 spp, depth 50, default camera of the generated scene; ids (xi=0.5 primary hits),
packed RGBA8, bit pattern of the FP32 mean radiance, ray count."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as O  # noqa: E402
from cudaraytracer_b200 import Scene, capi  # noqa: E402

SCENES = [("rtiow", capi.SCENE_RTIOW, 0), ("refdefault", capi.SCENE_REFERENCE_DEFAULT, 0), ("textured", capi.SCENE_TEXTURED, 256),
          ("stress", capi.SCENE_STRESS_10K, 20), ("grid", capi.SCENE_GRID_1M, 24)]

if __name__ == "__main__":
    W, H = 64, 36
    for name, kind, hint in SCENES:
        s = Scene.generate(kind, 1984, hint)
        cam = s.default_camera(W, H)
        b = O.refhost_render(s, cam, W, H, 2, 50, want_ids=True)
        np.savez_compressed(os.path.join(HERE, f"{name}_64x36_spp2.npz"), ids=b["ids"], rgba=b["rgba"],
                            mean_bits=b["mean"].view(np.uint32), rays=np.uint64(b["rays"]))
        print(name, s.n_hittables, "hit fraction", float((b["ids"] >= 0).mean()), "rays", b["rays"])
