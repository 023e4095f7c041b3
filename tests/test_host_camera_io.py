"""CPU tests of the host-side steps either side of the path: the scripted fly camera
(Renderer/Camera.cpp:37-118 restated in numpy here, float32 step by step) and the PPM reader/writer pair; and of the JPEG decoder
that replaces stbi_load (golden vectors from the reference's vendored decoder)."""
import ctypes as C
import os

import numpy as np
import pytest

from cudaraytracer_b200 import capi

F = np.float32


class RefCamera:
    """Camera.cpp / Camera.h in numpy float32 (the reference uses glm::vec3 floats)."""

    def __init__(self, pos):
        self.p = np.array(pos, dtype=F); self.o = np.array([0, 0, -1], dtype=F); self.u = np.array([0, 1, 0], dtype=F)
        self.speed = F(0.05); self.sens = F(0.1); self.yaw = F(270.0); self.pitch = F(0.0); self.fov = F(45.0)

    def inputs(self, keys, dx, dy, scroll):
        r = np.cross(self.o, self.u).astype(F)
        r = (r * (F(1.0) / np.sqrt((r * r).sum(dtype=F)))).astype(F)
        sp = self.speed
        if keys & capi.KEY_W: self.p = (self.p + sp * self.o).astype(F)
        if keys & capi.KEY_A: self.p = (self.p + sp * -r).astype(F)
        if keys & capi.KEY_S: self.p = (self.p + sp * -self.o).astype(F)
        if keys & capi.KEY_D: self.p = (self.p + sp * r).astype(F)
        if keys & capi.KEY_SPACE: self.p = (self.p + sp * self.u).astype(F)
        if keys & capi.KEY_LCTRL: self.p = (self.p + sp * -self.u).astype(F)
        self.speed = F(0.1) if keys & capi.KEY_LSHIFT else F(0.05)
        if keys & capi.KEY_C: self.p = np.array([0, 2, 12], dtype=F)
        if dx or dy:
            self.yaw = F(self.yaw + F(dx) * self.sens); self.pitch = F(np.clip(self.pitch + F(dy) * self.sens, -89, 89))
            k = F(np.pi / 180)
            d = np.array([np.cos(k * self.yaw) * np.cos(k * self.pitch), np.sin(k * self.pitch), np.sin(k * self.yaw) * np.cos(k * self.pitch)], dtype=F)
            self.o = (d / np.sqrt((d * d).sum(dtype=F))).astype(F)
        if scroll:
            self.fov = F(np.clip(self.fov - F(scroll), 1, 120))


def test_flycam_follows_the_reference_camera():
    lib = capi.load()
    cam = capi.FlyCam()
    pos = (C.c_float * 3)(0.0, 2.0, 12.0)
    assert lib.crt_flycam_init(C.byref(cam), pos) == 0
    ref = RefCamera((0, 2, 12))
    rng = np.random.default_rng(3)
    moved = C.c_int()
    for step in range(400):
        keys = int(rng.integers(0, 128)) if step % 3 else 0
        if step == 250: keys |= capi.KEY_C
        dx, dy = (float(rng.integers(-30, 30)), float(rng.integers(-30, 30))) if step % 5 == 0 else (0.0, 0.0)
        scroll = float(rng.integers(-3, 4)) if step % 7 == 0 else 0.0
        before = (tuple(cam.position), tuple(cam.orientation), cam.fov_degrees)
        assert lib.crt_flycam_step(C.byref(cam), keys, dx, dy, scroll, C.byref(moved)) == 0
        ref.inputs(keys, dx, dy, scroll)
        assert np.allclose(np.array(cam.position[:]), ref.p, rtol=0, atol=2e-5), step
        assert np.allclose(np.array(cam.orientation[:]), ref.o, rtol=0, atol=2e-6), step
        assert abs(cam.fov_degrees - float(ref.fov)) == 0 and abs(cam.pitch) <= 89.0
        assert bool(moved.value) == (before != (tuple(cam.position), tuple(cam.orientation), cam.fov_degrees))
    # the mapping to InputStruct is CudaLayer.cpp:413-431: upV = normalize(cross(orientation, normalize(cross(orientation, up))))
    out = capi.InputStruct()
    assert lib.crt_flycam_inputs(C.byref(cam), None, None, C.byref(out)) == 0
    r = np.cross(ref.o, ref.u); r /= np.linalg.norm(r)
    u = np.cross(ref.o, r); u /= np.linalg.norm(u)
    assert np.allclose([out.up_x, out.up_y, out.up_z], u, atol=1e-5)
    assert np.isclose(out.fov, np.radians(float(ref.fov)), rtol=1e-6) and (out.near_plane, out.far_plane) == (F(0.1), F(10.0))


def test_ppm_write_then_load_round_trip(tmp_path):
    lib = capi.load()
    W, H = 37, 19
    rng = np.random.default_rng(5)
    rgb = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    packed = (rgb[..., 0].astype(np.uint32) | (rgb[..., 1].astype(np.uint32) << 8) | (rgb[..., 2].astype(np.uint32) << 16) | np.uint32(0xFF000000))
    packed = np.ascontiguousarray(packed)
    path = str(tmp_path / "t.ppm").encode()
    assert lib.crt_write_ppm(path, packed.ctypes.data_as(C.c_void_p), W, H, 0) == 0
    px = C.POINTER(C.c_uint8)(); w = C.c_int32(); h = C.c_int32()
    assert lib.crt_load_ppm(path, C.byref(px), C.byref(w), C.byref(h)) == 0 and (w.value, h.value) == (W, H)
    got = np.ctypeslib.as_array(px, shape=(H, W, 3)).copy()
    lib.crt_free(px)
    assert np.array_equal(got, rgb)
    # a header with a comment, and the error paths
    open(tmp_path / "c.ppm", "wb").write(b"P6\n# made by hand\n2 1\n255\n" + bytes([1, 2, 3, 4, 5, 6]))
    assert lib.crt_load_ppm(str(tmp_path / "c.ppm").encode(), C.byref(px), C.byref(w), C.byref(h)) == 0 and (w.value, h.value) == (2, 1)
    assert list(np.ctypeslib.as_array(px, shape=(6,))) == [1, 2, 3, 4, 5, 6]
    lib.crt_free(px)
    open(tmp_path / "bad.ppm", "wb").write(b"P5\n2 1\n255\n..")
    assert lib.crt_load_ppm(str(tmp_path / "bad.ppm").encode(), C.byref(px), C.byref(w), C.byref(h)) != 0
    open(tmp_path / "short.ppm", "wb").write(b"P6\n4 4\n255\n....")
    assert lib.crt_load_ppm(str(tmp_path / "short.ppm").encode(), C.byref(px), C.byref(w), C.byref(h)) != 0
    assert lib.crt_load_ppm(b"/nonexistent/x.ppm", C.byref(px), C.byref(w), C.byref(h)) != 0


# ---- JPEG decode (csrc/jpeg_decode.cpp replaces stbi_load of Utils/RawStbImage.h:11-22) ----
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
STB_REF_SO = os.path.join(os.path.dirname(os.path.dirname(GOLDEN)), "oracle", "_ref", "libstb_ref.so")
REF_TEXTURES = "/root/reference/assets/textures"


def _decode_jpeg(blob):
    lib = capi.load()
    px = C.POINTER(C.c_uint8)()
    w, h, n = C.c_int32(), C.c_int32(), C.c_int32()
    rc = lib.crt_decode_jpeg(blob, len(blob), C.byref(px), C.byref(w), C.byref(h), C.byref(n))
    if rc != 0:
        return rc, None, 0
    a = np.ctypeslib.as_array(px, shape=(h.value, w.value, 3)).copy()
    lib.crt_free(px)
    return 0, a, n.value


def _stb_decode(blob):
    stb = C.CDLL(STB_REF_SO)
    stb.stbi_load_from_memory.restype = C.POINTER(C.c_uint8)
    stb.stbi_load_from_memory.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    stb.stbi_image_free.argtypes = [C.c_void_p]
    w, h, n = C.c_int(), C.c_int(), C.c_int()
    p = stb.stbi_load_from_memory(blob, len(blob), C.byref(w), C.byref(h), C.byref(n), 3)
    if not p:
        return None, 0
    a = np.ctypeslib.as_array(p, shape=(h.value, w.value, 3)).copy()
    stb.stbi_image_free(p)
    return a, n.value


def test_jpeg_golden_vectors_bit_exact():
    """Every committed fixture decodes to exactly the RGB the reference's vendored decoder produced for it
    (tests/golden/make_jpeg_fixtures.py): baseline 4:4:4 / 4:2:2 / 4:2:0, progressive, restart intervals, grey, 1x1."""
    exp = np.load(os.path.join(GOLDEN, "jpeg_expected.npz"))
    names = sorted(k for k in exp.files if not k.endswith("_n"))
    assert len(names) == 7
    for name in names:
        blob = open(os.path.join(GOLDEN, "jpeg", name + ".jpg"), "rb").read()
        rc, got, n = _decode_jpeg(blob)
        assert rc == 0, name
        assert n == int(exp[name + "_n"]), name
        assert got.shape == exp[name].shape and np.array_equal(got, exp[name]), name


def test_jpeg_thread_count_does_not_change_the_result(monkeypatch):
    blob = open(os.path.join(GOLDEN, "jpeg", "base444_rst.jpg"), "rb").read()
    monkeypatch.setenv("CRT_JPEG_THREADS", "1")
    _, a, _ = _decode_jpeg(blob)
    monkeypatch.setenv("CRT_JPEG_THREADS", "7")
    _, b, _ = _decode_jpeg(blob)
    assert np.array_equal(a, b)


def test_jpeg_errors_and_load_image(tmp_path):
    lib = capi.load()
    good = open(os.path.join(GOLDEN, "jpeg", "base420.jpg"), "rb").read()
    assert _decode_jpeg(b"\x89PNG\r\n\x1a\n" + b"\0" * 32)[0] == 1            # not a JPEG: CRT_ERR_INVALID
    assert _decode_jpeg(good[:40])[0] == 1                                      # cut inside the headers
    assert _decode_jpeg(good[: good.index(b"\xff\xda")])[0] == 1                # no scan at all
    arith = good.replace(b"\xff\xc0", b"\xff\xc9", 1)                           # SOF9: arithmetic coding
    assert _decode_jpeg(arith)[0] == 4                                          # CRT_ERR_UNSUPPORTED
    rc, cut, _ = _decode_jpeg(good[: len(good) - len(good) // 3])               # truncated entropy data: what was decoded is kept
    assert rc == 0 and cut.shape == (38, 50, 3)
    # crt_load_image: JPEG and PPM by magic number, anything else refused
    exp = np.load(os.path.join(GOLDEN, "jpeg_expected.npz"))["base420"]
    px = C.POINTER(C.c_uint8)()
    w, h, n = C.c_int32(), C.c_int32(), C.c_int32()
    path = os.path.join(GOLDEN, "jpeg", "base420.jpg").encode()
    assert lib.crt_load_image(path, C.byref(px), C.byref(w), C.byref(h), C.byref(n)) == 0
    assert (w.value, h.value, n.value) == (50, 38, 3)
    assert np.array_equal(np.ctypeslib.as_array(px, shape=(38, 50, 3)), exp)
    lib.crt_free(px)
    (tmp_path / "t.ppm").write_bytes(b"P6\n2 1\n255\n" + bytes([1, 2, 3, 4, 5, 6]))
    assert lib.crt_load_image(str(tmp_path / "t.ppm").encode(), C.byref(px), C.byref(w), C.byref(h), C.byref(n)) == 0
    assert (w.value, h.value, n.value) == (2, 1, 3) and list(np.ctypeslib.as_array(px, shape=(6,))) == [1, 2, 3, 4, 5, 6]
    lib.crt_free(px)
    (tmp_path / "junk.bin").write_bytes(b"GIF89a" + b"\0" * 20)
    assert lib.crt_load_image(str(tmp_path / "junk.bin").encode(), C.byref(px), C.byref(w), C.byref(h), C.byref(n)) == 4
    assert lib.crt_load_image(b"/nonexistent/x.jpg", C.byref(px), C.byref(w), C.byref(h), C.byref(n)) == 1


@pytest.mark.skipif(not os.path.exists(STB_REF_SO), reason="oracle/_ref/libstb_ref.so not built (needs /root/reference)")
def test_jpeg_matches_vendored_decoder_on_generated_files():
    """Fresh files (Pillow encodes; sizes around the MCU boundaries, every chroma layout, both processes, restart intervals)
    against the decoder the reference vendors, compiled from where it lies."""
    Image = pytest.importorskip("PIL.Image")
    import io
    rng = np.random.default_rng(11)
    n_checked = 0
    for (w, h) in [(1, 1), (2, 3), (8, 8), (9, 7), (15, 16), (16, 17), (31, 33), (47, 19), (100, 75)]:
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        img[: h // 2] = (img[: h // 2].astype(np.int32) // 8 + np.arange(w)[None, :, None] * 200 // max(w, 1)).clip(0, 255).astype(np.uint8)
        for sub in (0, 1, 2):
            for prog in (False, True):
                for extra in ({}, {"restart_marker_blocks": 2}):
                    bio = io.BytesIO()
                    Image.fromarray(img).save(bio, "JPEG", quality=int(rng.integers(20, 96)), subsampling=sub, progressive=prog, **extra)
                    blob = bio.getvalue()
                    want, wn = _stb_decode(blob)
                    rc, got, n = _decode_jpeg(blob)
                    assert want is not None and rc == 0
                    assert n == wn and np.array_equal(got, want), (w, h, sub, prog, extra)
                    n_checked += 1
        bio = io.BytesIO()
        Image.fromarray(img[..., 1]).save(bio, "JPEG", quality=70)
        want, wn = _stb_decode(bio.getvalue())
        rc, got, n = _decode_jpeg(bio.getvalue())
        assert rc == 0 and n == wn == 1 and np.array_equal(got, want)
    assert n_checked == 9 * 12


@pytest.mark.skipif(not (os.path.exists(STB_REF_SO) and os.path.isdir(REF_TEXTURES)),
                    reason="needs the reference's texture assets and the vendored decoder (build container only)")
def test_jpeg_reference_texture_assets_bit_exact():
    """The textures the reference ships (8192x4096 baseline 4:4:4 with restart intervals; a progressive 4:2:0 logo): the
    texels this library hands to Image::value are the ones stbi_load gives the reference."""
    import glob
    files = sorted(glob.glob(os.path.join(REF_TEXTURES, "*.jpg")))
    assert files
    for f in files:
        if os.path.getsize(f) > 2_500_000 and not f.endswith("8k_earth_nightmap.jpg"):
            continue  # one of the 8k files is enough for the CPU suite's time budget
        blob = open(f, "rb").read()
        want, wn = _stb_decode(blob)
        rc, got, n = _decode_jpeg(blob)
        assert rc == 0 and n == wn and np.array_equal(got, want), f


@pytest.mark.skipif(not (os.path.exists(STB_REF_SO) and os.path.isdir(REF_TEXTURES)),
                    reason="needs the reference's texture assets and the vendored decoder (build container only)")
def test_textured_scene_takes_the_reference_assets_when_pointed_at_them(monkeypatch):
    """C3's earth is assets/textures/8k_earth_nightmap.jpg and its small spheres UTH-logo-english.jpg. With
    CRT_ASSET_DIR set the generator decodes them (same texels as stbi_load); without it, procedural stand-ins of equal size."""
    from cudaraytracer_b200 import Scene, load_image
    monkeypatch.setenv("CRT_ASSET_DIR", REF_TEXTURES)
    s = Scene.generate(capi.SCENE_TEXTURED, 1984, 0)
    d = s.desc
    assert d.n_images == 2 and (d.images[0].width, d.images[0].height) == (8192, 4096) and (d.images[1].width, d.images[1].height) == (1241, 1241)
    logo = np.ctypeslib.as_array(d.images[1].pixels, shape=(1241, 1241, 3))
    want, _ = _stb_decode(open(os.path.join(REF_TEXTURES, "UTH-logo-english.jpg"), "rb").read())
    assert np.array_equal(logo, want)
    assert np.array_equal(load_image(os.path.join(REF_TEXTURES, "UTH-logo-english.jpg")), want)
    monkeypatch.delenv("CRT_ASSET_DIR")
    s2 = Scene.generate(capi.SCENE_TEXTURED, 1984, 0)  # the Scene owns the arrays its desc points into
    p = s2.desc
    assert (p.images[0].width, p.images[0].height) == (8192, 4096)
    assert not np.array_equal(np.ctypeslib.as_array(p.images[1].pixels, shape=(1241, 1241, 3)), want)


def test_png_writer_round_trip(tmp_path):
    """crt_write_png: chunk CRCs, a zlib stream every inflater accepts, the pixels of the frame (rows flipped on request)."""
    import struct
    import zlib
    from cudaraytracer_b200 import write_png
    rng = np.random.default_rng(5)
    for (W, H) in [(1, 1), (37, 23), (300, 250)]:  # the last one needs more than one stored block (225 750 raw bytes)
        rgb = rng.integers(0, 256, (H, W, 3), dtype=np.uint32)
        img = rgb[..., 0] | rgb[..., 1] << 8 | rgb[..., 2] << 16 | np.uint32(0xFF000000)
        for flip in (False, True):
            path = str(tmp_path / f"f_{W}_{int(flip)}.png")
            write_png(path, img, flip_rows=flip)
            data = open(path, "rb").read()
            assert data[:8] == b"\x89PNG\r\n\x1a\n"
            pos, chunks = 8, []
            while pos < len(data):
                n, typ = struct.unpack(">I4s", data[pos:pos + 8])
                body = data[pos + 8:pos + 8 + n]
                (crc,) = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])
                assert crc == zlib.crc32(typ + body) & 0xFFFFFFFF
                chunks.append((typ, body))
                pos += 12 + n
            assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
            assert struct.unpack(">IIBBBBB", chunks[0][1]) == (W, H, 8, 2, 0, 0, 0)
            raw = np.frombuffer(zlib.decompress(chunks[1][1]), dtype=np.uint8).reshape(H, 1 + 3 * W)
            assert (raw[:, 0] == 0).all()
            want = rgb[::-1] if flip else rgb
            assert np.array_equal(raw[:, 1:].reshape(H, W, 3), want.astype(np.uint8))
            try:
                from PIL import Image
                assert np.array_equal(np.asarray(Image.open(path).convert("RGB")), want.astype(np.uint8))
            except ImportError:
                pass
    assert capi.load().crt_write_png(b"/nonexistent/dir/x.png", img.ctypes.data_as(C.c_void_p), 300, 250, 0) != 0
