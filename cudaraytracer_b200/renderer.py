"""Thin Python host over the C ABI (include/crt.h) for tests and bench.py.

Names follow the reference's domain: a Scene is the hittable list + materials + textures
(CudaLayer::GenerateWorld, CudaLayer.cpp:103-256), the camera is its InputStruct
(SharedStructs.h:3-24), Renderer.render() is LaunchKernel (Kernel.cu:178-191) with a progressive
accumulation buffer. All rendering happens in libcrt.so on the GPU.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import capi


class CrtError(RuntimeError):
    pass


def _check(status: int, ctx=None):
    if status != capi.CRT_OK:
        lib = capi.load()
        msg = lib.crt_last_error_string(ctx)
        raise CrtError(f"crt status {status}: {msg.decode() if msg else '?'}")


class Scene:
    """Owns a crt_scene_builder (generated scene) or numpy-backed arrays (custom scene)."""

    def __init__(self):
        self._builder = None
        self._keep = []
        self._lib = None
        self.desc = None

    @classmethod
    def generate(cls, kind: int, seed: int = 1984, size_hint: int = 0, host_only: bool = False) -> "Scene":
        """host_only: use libcrt_scenegen.so (no CUDA, the product library stays unmapped) instead of libcrt.so."""
        lib = capi.load_scenegen() if host_only else capi.load()
        s = cls()
        s._lib = lib
        b = C.c_void_p()
        st = lib.crt_scene_generate(kind, seed, size_hint, C.byref(b))
        if st != capi.CRT_OK:
            msg = lib.crt_last_error_string(None)
            raise CrtError(f"crt status {st}: {msg.decode() if msg else '?'}")
        s._builder = b
        s.desc = lib.crt_scene_builder_desc(b).contents
        return s

    @classmethod
    def from_arrays(cls, hittables, materials, textures, images=()) -> "Scene":
        """hittables: list of dict(type, cx, cy, cz, a, b=0, material, active=1); materials: dict(type, texture, param,
        light_intensity); textures: dict(type, image, color, odd, even); images: list of HxWx3 uint8 arrays or None."""
        s = cls()
        H = (capi.Hittable * max(len(hittables), 1))()
        for i, h in enumerate(hittables):
            H[i] = capi.Hittable(h["type"], h.get("active", 1), h["cx"], h["cy"], h["cz"], h["a"], h.get("b", 0.0), h["material"])
        M = (capi.Material * max(len(materials), 1))()
        for i, m in enumerate(materials):
            M[i] = capi.Material(m["type"], m.get("texture", -1), m.get("param", 0.0), m.get("light_intensity", 0))
        T = (capi.Texture * max(len(textures), 1))()
        for i, t in enumerate(textures):
            T[i] = capi.Texture(t["type"], t.get("image", -1), (C.c_float * 3)(*t.get("color", (0, 0, 0))),
                                (C.c_float * 3)(*t.get("odd", (0, 0, 0))), (C.c_float * 3)(*t.get("even", (0, 0, 0))), 0)
        I = (capi.Image * max(len(images), 1))()
        for i, im in enumerate(images):
            if im is None:
                I[i] = capi.Image(None, 0, 0)
            else:
                arr = np.ascontiguousarray(im, dtype=np.uint8)
                assert arr.ndim == 3 and arr.shape[2] == 3
                s._keep.append(arr)
                I[i] = capi.Image(arr.ctypes.data_as(C.POINTER(C.c_uint8)), arr.shape[1], arr.shape[0])
        s._keep += [H, M, T, I]
        s.desc = capi.SceneDesc(H, len(hittables), M, len(materials), T, len(textures), I, len(images))
        return s

    def default_camera(self, width: int, height: int) -> capi.InputStruct:
        if self._builder is None:
            raise CrtError("default_camera: only generated scenes carry a camera")
        cam = capi.InputStruct()
        if self._lib.crt_scene_builder_camera(self._builder, width, height, C.byref(cam)) != capi.CRT_OK:
            raise CrtError("crt_scene_builder_camera failed")
        return cam

    @property
    def n_hittables(self) -> int:
        return int(self.desc.n_hittables)

    def hittable_array(self) -> np.ndarray:
        """(n, 8) view: type, active, cx, cy, cz, a, b, material (floats reinterpreted per column by the caller)."""
        n = self.n_hittables
        buf = C.string_at(self.desc.hittables, n * C.sizeof(capi.Hittable))
        return np.frombuffer(buf, dtype=np.dtype([("type", "<i4"), ("active", "<i4"), ("cx", "<f4"), ("cy", "<f4"), ("cz", "<f4"),
                                                  ("a", "<f4"), ("b", "<f4"), ("material", "<i4")]))

    def host_bytes(self) -> int:
        d = self.desc
        n = d.n_hittables * 32 + d.n_materials * 16 + d.n_textures * 48
        for i in range(d.n_images):
            if d.images[i].pixels:
                n += d.images[i].width * d.images[i].height * 3
        return int(n)

    def __del__(self):
        if self._builder is not None:
            try:
                self._lib.crt_scene_builder_free(self._builder)
            except Exception:
                pass
            self._builder = None


def load_image(path: str) -> np.ndarray:
    """JPEG or binary PPM file -> (H, W, 3) uint8, the array Scene.from_arrays takes as an image texture.
    The library's own decoder (csrc/jpeg_decode.cpp) in place of the reference's stbi_load (Utils/RawStbImage.h:11-22)."""
    lib = capi.load()
    px = C.POINTER(C.c_uint8)()
    w, h, n = C.c_int32(), C.c_int32(), C.c_int32()
    _check(lib.crt_load_image(os.fsencode(path), C.byref(px), C.byref(w), C.byref(h), C.byref(n)))
    try:
        return np.ctypeslib.as_array(px, shape=(h.value, w.value, 3)).copy()
    finally:
        lib.crt_free(px)


def look_camera(position, orientation, fov_deg=45.0, near=0.1, far=10.0, up=(0.0, 1.0, 0.0),
                bg0=(1.0, 1.0, 1.0), bg1=(0.5, 0.7, 1.0)) -> capi.InputStruct:
    """Camera fields -> InputStruct exactly as CudaLayer::OnAttach does (CudaLayer.cpp:45-62)."""
    f3 = C.c_float * 3
    cam = capi.InputStruct()
    _check(capi.load().crt_camera_look(f3(*position), f3(*orientation), f3(*up), fov_deg, near, far, f3(*bg0), f3(*bg1), C.byref(cam)))
    return cam


class Renderer:
    def __init__(self, device: int = 0):
        self.lib = capi.load()
        self.ctx = C.c_void_p()
        _check(self.lib.crt_context_create(device, C.byref(self.ctx)))
        self.width = self.height = 0

    def close(self):
        if self.ctx:
            self.lib.crt_context_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    stream_handle = None

    def _ck(self, st):
        _check(st, self.ctx)

    def set_stream(self, cuda_stream: int):
        self._ck(self.lib.crt_set_stream(self.ctx, C.c_void_p(cuda_stream)))
        self.stream_handle = int(cuda_stream)  # None until set: the context then runs on its own non-blocking stream

    def set_bvh_builder(self, builder: int):
        """capi.BUILDER_AUTO / BUILDER_HOST_SAH (threaded SAH on the host) / BUILDER_GPU_LBVH (Morton LBVH on the device)."""
        self._ck(self.lib.crt_set_bvh_builder(self.ctx, builder))

    def set_scene_cache(self, enabled: bool):
        """crt_scene_upload of a description whose content hash equals the scene already on the device returns at once (default
        on); off = every upload validates, builds and copies."""
        self._ck(self.lib.crt_set_scene_cache(self.ctx, int(bool(enabled))))

    def upload(self, scene: Scene):
        self._ck(self.lib.crt_scene_upload(self.ctx, C.byref(scene.desc)))

    def edit_hittable(self, index: int, **fields):
        """Stage an edit of hittable `index` (fields of crt_hittable: type, active, cx, cy, cz, a, b, material)."""
        self._ck(self.lib.crt_scene_edit_hittable(self.ctx, index, C.byref(capi.Hittable(**fields))))

    def edit_material(self, index: int, **fields):
        self._ck(self.lib.crt_scene_edit_material(self.ctx, index, C.byref(capi.Material(**fields))))

    def commit(self, force_rebuild: bool = False) -> bool:
        """Apply the staged edits: GPU refit, or a rebuild (host SAH / GPU LBVH) when the topology changed. Returns True if rebuilt."""
        rebuilt = C.c_int()
        self._ck(self.lib.crt_scene_commit(self.ctx, int(force_rebuild), C.byref(rebuilt)))
        return bool(rebuilt.value)

    def bvh_stats(self) -> capi.BvhStats:
        st = capi.BvhStats()
        self._ck(self.lib.crt_bvh_get_stats(self.ctx, C.byref(st)))
        return st

    def set_camera(self, cam: capi.InputStruct):
        self._ck(self.lib.crt_set_camera(self.ctx, C.byref(cam)))

    def set_lens(self, aperture: float, focus_dist: float):
        self._ck(self.lib.crt_set_lens(self.ctx, aperture, focus_dist))

    def resize(self, width: int, height: int):
        self._ck(self.lib.crt_framebuffer_resize(self.ctx, width, height))
        self.width, self.height = width, height

    def clear(self):
        self._ck(self.lib.crt_accum_clear(self.ctx))

    def set_accum_device_buffer(self, ptr: int):
        self._ck(self.lib.crt_set_accum_device_buffer(self.ctx, C.c_void_p(ptr)))

    def render(self, spp: int, max_depth: int, spp_begin: int = 0, spp_stride: int = 1, seed: int = 1984,
               variant: int = capi.VARIANT_AUTO, jitter: int = capi.JITTER_RANDOM, flags: int = 0, row_tiles: int = 0):
        """row_tiles = capi.row_tiles(rank, world, rows_per_tile): only that row-tile shard of the frame (0 = every row)."""
        p = capi.RenderParams(spp_begin, spp, spp_stride, max_depth, seed, variant, jitter, flags, row_tiles)
        self._ck(self.lib.crt_render(self.ctx, C.byref(p)))

    def synchronize(self):
        self._ck(self.lib.crt_synchronize(self.ctx))

    def last_render_ms(self) -> float:
        ms = C.c_float()
        self._ck(self.lib.crt_last_render_ms(self.ctx, C.byref(ms)))
        return float(ms.value)

    def last_variant(self) -> int:
        """The variant the last render() ran (capi.VARIANT_*): what AUTO resolved to, measured where both kernels apply."""
        return int(self.lib.crt_last_variant(self.ctx))

    def resolve(self, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.height, self.width), dtype=np.uint32)
        assert out.dtype == np.uint32 and out.size == self.width * self.height and out.flags.c_contiguous
        self._ck(self.lib.crt_resolve(self.ctx, out.ctypes.data_as(C.c_void_p)))
        return out

    def resolve_device(self, ptr: int):
        self._ck(self.lib.crt_resolve_device(self.ctx, C.c_void_p(ptr)))

    def resolve_rows(self, row_tiles: int, out_rgba_ptr: int):
        """Resolve the rows of a row-tile shard from this context's buffer into the frame at out_rgba_ptr (may be a peer mapping)."""
        self._ck(self.lib.crt_resolve_rows(self.ctx, row_tiles, C.c_void_p(out_rgba_ptr)))

    def read_rgba(self, out: np.ndarray | None = None) -> np.ndarray:
        """The context's own RGBA8 frame as it stands (written by resolve() or by the peers' resolve_peers)."""
        if out is None:
            out = np.empty((self.height, self.width), dtype=np.uint32)
        assert out.dtype == np.uint32 and out.size == self.width * self.height and out.flags.c_contiguous
        self._ck(self.lib.crt_read_rgba(self.ctx, out.ctypes.data_as(C.c_void_p)))
        return out

    def accum_ptr(self) -> int:
        return int(self.lib.crt_accum_device_ptr(self.ctx) or 0)

    def rgba_ptr(self) -> int:
        return int(self.lib.crt_rgba_device_ptr(self.ctx) or 0)

    def ipc_export(self, which: int) -> bytes:
        """64-byte CUDA IPC handle of the context's own accumulation buffer (capi.BUFFER_ACCUM) or RGBA8 frame (BUFFER_RGBA)."""
        buf = (C.c_uint8 * capi.IPC_HANDLE_BYTES)()
        self._ck(self.lib.crt_ipc_export(self.ctx, which, buf))
        return bytes(buf)

    def ipc_open(self, handle: bytes) -> int:
        """Map a buffer another PROCESS exported; returns the device pointer valid on this context's device."""
        assert len(handle) == capi.IPC_HANDLE_BYTES
        buf = (C.c_uint8 * capi.IPC_HANDLE_BYTES).from_buffer_copy(handle)
        out = C.c_void_p()
        self._ck(self.lib.crt_ipc_open(self.ctx, buf, C.byref(out)))
        return int(out.value)

    def ipc_close(self, ptr: int):
        self._ck(self.lib.crt_ipc_close(self.ctx, C.c_void_p(ptr)))

    def resolve_peers(self, peer_ptrs, row_begin: int, row_end: int, out_rgba_ptr: int, out_sum_ptr: int = 0):
        """Sum the accumulation buffers `peer_ptrs` over rows [row_begin, row_end), resolve, store into the frame at out_rgba_ptr."""
        arr = (C.c_void_p * len(peer_ptrs))(*[C.c_void_p(int(p)) for p in peer_ptrs])
        self._ck(self.lib.crt_resolve_peers(self.ctx, arr, len(peer_ptrs), row_begin, row_end, C.c_void_p(out_rgba_ptr),
                                            C.c_void_p(out_sum_ptr) if out_sum_ptr else None))

    def read_accum(self) -> np.ndarray:
        out = np.empty((self.height, self.width, 4), dtype=np.float32)
        self._ck(self.lib.crt_read_accum(self.ctx, out.ctypes.data_as(C.c_void_p)))
        return out

    def read_accum_static(self, device_ptr: int = 0, n_pixels: int = None) -> np.ndarray:
        """crt_read_accum_static: the float4 view of n_pixels entries of an int64x4 accumulation buffer (default: this renderer's
        own buffer, all of it) through the library's STATIC staging array - no allocation on the device side of the call."""
        n = self.width * self.height if n_pixels is None else int(n_pixels)
        out = np.empty((n, 4), dtype=np.float32)
        ptr = device_ptr or self.accum_ptr()
        self._ck(self.lib.crt_read_accum_static(self.ctx, C.c_void_p(ptr), n, out.ctypes.data_as(C.c_void_p)))
        return out

    def primary_ids(self, jitter: int = capi.JITTER_CENTER, seed: int = 1984) -> np.ndarray:
        out = np.empty((self.height, self.width), dtype=np.int32)
        self._ck(self.lib.crt_primary_ids(self.ctx, jitter, seed, out.ctypes.data_as(C.c_void_p)))
        return out

    def counters(self) -> capi.Counters:
        c = capi.Counters()
        self._ck(self.lib.crt_get_counters(self.ctx, C.byref(c)))
        return c

    def reset_counters(self):
        self._ck(self.lib.crt_reset_counters(self.ctx))


def unpack_rgba8(img: np.ndarray) -> np.ndarray:
    """(H, W) packed A<<24|B<<16|G<<8|R -> (H, W, 3) uint8 RGB."""
    return np.stack([img & 255, (img >> 8) & 255, (img >> 16) & 255], axis=-1).astype(np.uint8)


def write_ppm(path: str, img: np.ndarray, flip_rows: bool = True):
    img = np.ascontiguousarray(img, dtype=np.uint32)
    _check(capi.load().crt_write_ppm(path.encode(), img.ctypes.data_as(C.c_void_p), img.shape[1], img.shape[0], int(flip_rows)))


def write_png(path: str, img: np.ndarray, flip_rows: bool = True):
    img = np.ascontiguousarray(img, dtype=np.uint32)
    _check(capi.load().crt_write_png(path.encode(), img.ctypes.data_as(C.c_void_p), img.shape[1], img.shape[0], int(flip_rows)))
