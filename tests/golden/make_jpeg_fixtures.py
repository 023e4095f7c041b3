#!/usr/bin/env python
"""Generates tests/golden/jpeg/*.jpg and tests/golden/jpeg_expected.npz (run in the build container, where /root/reference is
mounted and oracle/_ref/libstb_ref.so has been built from the reference's vendored vendor/stb/stb_image.h by `make -C oracle`).

The JPEG files are synthetic (Pillow encodes a deterministic pattern); the expected RGB arrays are what the VENDORED decoder
(stbi_load_from_memory, req_comp = 3) returns for each file: the golden vectors that pin cudaraytracer_b200/csrc/jpeg_decode.cpp
on machines where the reference tree is absent."""
import ctypes as C
import io
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def stb_decode(blob):
    stb = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libstb_ref.so"))
    stb.stbi_load_from_memory.restype = C.POINTER(C.c_uint8)
    stb.stbi_load_from_memory.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    stb.stbi_image_free.argtypes = [C.c_void_p]
    w, h, n = C.c_int(), C.c_int(), C.c_int()
    p = stb.stbi_load_from_memory(blob, len(blob), C.byref(w), C.byref(h), C.byref(n), 3)
    assert p, "vendored decoder rejected the file"
    a = np.ctypeslib.as_array(p, shape=(h.value, w.value, 3)).copy()
    stb.stbi_image_free(p)
    return a, n.value


def pattern(w, h, seed):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    img = np.stack([x * 255 // max(w - 1, 1), y * 255 // max(h - 1, 1), (x * 5 + y * 3) % 256], -1).astype(np.int32)
    img += rng.integers(-24, 24, img.shape)
    img[h // 3: h // 2, w // 4: w // 2] = (250, 10, 30)  # a hard edge: exercises the chroma filters and the clamps
    return img.clip(0, 255).astype(np.uint8)


CASES = {
    # name: (width, height, Pillow save options, grey)
    "base444_rst": (67, 41, dict(quality=88, subsampling=0, restart_marker_blocks=3), False),
    "base420": (50, 38, dict(quality=75, subsampling=2), False),
    "base422_q30": (33, 17, dict(quality=30, subsampling=1), False),
    "prog420": (61, 45, dict(quality=80, subsampling=2, progressive=True), False),
    "prog444_rst": (40, 40, dict(quality=92, subsampling=0, progressive=True, restart_marker_rows=1), False),
    "grey": (29, 23, dict(quality=85), True),
    "tiny": (1, 1, dict(quality=90, subsampling=2), False),
}

if __name__ == "__main__":
    os.makedirs(os.path.join(HERE, "jpeg"), exist_ok=True)
    expected = {}
    for i, (name, (w, h, opts, grey)) in enumerate(CASES.items()):
        img = pattern(w, h, 100 + i)
        im = Image.fromarray(img[..., 0] if grey else img)
        bio = io.BytesIO()
        im.save(bio, "JPEG", **opts)
        blob = bio.getvalue()
        open(os.path.join(HERE, "jpeg", name + ".jpg"), "wb").write(blob)
        rgb, n = stb_decode(blob)
        expected[name] = rgb
        expected[name + "_n"] = np.int32(n)
        print(name, rgb.shape, n, len(blob), "bytes")
    np.savez_compressed(os.path.join(HERE, "jpeg_expected.npz"), **expected)
