"""CPU tests of the host-side steps either side of the path: the scripted fly camera
(Renderer/Camera.cpp:37-118 restated in numpy here, float32 step by step) and the PPM reader/writer pair."""
import ctypes as C

import numpy as np

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
