"""ctypes binding of include/crt.h (cudaraytracer_b200/libcrt.so).

This is plumbing for tests/ and bench.py: the product is the C-ABI library itself. Loading fails loudly
if the library has not been built (``python -c 'import __graft_entry__ as g; g.build()'`` or
``make -C cudaraytracer_b200/csrc``); nothing here falls back to a CPU implementation.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CRT_LIB_PATH") or os.path.join(_HERE, "libcrt.so")  # override: A/B experiments only

# enums of include/crt.h
CRT_OK, CRT_ERR_INVALID, CRT_ERR_CUDA, CRT_ERR_NOMEM, CRT_ERR_UNSUPPORTED, CRT_ERR_STATE = 0, 1, 2, 3, 4, 5
SPHERE, XYRECT, XZRECT, YZRECT = 0, 1, 2, 3
LAMBERTIAN, METAL, DIELECTRIC, DIFFUSELIGHT = 0, 1, 2, 3
CONSTANT, CHECKER, IMAGE = 0, 1, 2
STATIC_ACCUM_MAX_PIXELS = 262144  # CRT_STATIC_ACCUM_MAX_PIXELS, include/crt.h
BUILDER_AUTO, BUILDER_HOST_SAH, BUILDER_GPU_LBVH = 0, 1, 2
VARIANT_AUTO, VARIANT_MEGAKERNEL, VARIANT_WAVEFRONT, VARIANT_PIPELINE = 0, 1, 2, 3
JITTER_RANDOM, JITTER_CENTER = 0, 1
FLAG_EXACT_MATH, FLAG_COUNT_TESTS, FLAG_TRUNCATED_GRID = 1, 2, 4
BUFFER_ACCUM, BUFFER_RGBA = 0, 1


def row_tiles(rank: int, world: int, rows_per_tile: int = 16) -> int:
    """CRT_ROW_TILES of include/crt.h."""
    if not (0 <= rank < world <= 255 and 1 <= rows_per_tile <= 65535):
        raise ValueError("row_tiles: need 0 <= rank < world <= 255 and 1 <= rows_per_tile <= 65535")
    return (rows_per_tile << 16) | (world << 8) | rank

IPC_HANDLE_BYTES = 64
SCENE_RTIOW, SCENE_TEXTURED, SCENE_GRID_1M, SCENE_STRESS_10K, SCENE_REFERENCE_DEFAULT = 1, 2, 3, 4, 5


class Hittable(C.Structure):
    _fields_ = [("type", C.c_int32), ("active", C.c_int32), ("cx", C.c_float), ("cy", C.c_float), ("cz", C.c_float),
                ("a", C.c_float), ("b", C.c_float), ("material", C.c_int32)]


class Material(C.Structure):
    _fields_ = [("type", C.c_int32), ("texture", C.c_int32), ("param", C.c_float), ("light_intensity", C.c_int32)]


class Texture(C.Structure):
    _fields_ = [("type", C.c_int32), ("image", C.c_int32), ("color", C.c_float * 3), ("odd", C.c_float * 3),
                ("even", C.c_float * 3), ("pad_", C.c_int32)]


class Image(C.Structure):
    _fields_ = [("pixels", C.POINTER(C.c_uint8)), ("width", C.c_int32), ("height", C.c_int32)]


class SceneDesc(C.Structure):
    _fields_ = [("hittables", C.POINTER(Hittable)), ("n_hittables", C.c_uint64),
                ("materials", C.POINTER(Material)), ("n_materials", C.c_uint64),
                ("textures", C.POINTER(Texture)), ("n_textures", C.c_uint64),
                ("images", C.POINTER(Image)), ("n_images", C.c_uint64)]


class InputStruct(C.Structure):
    _fields_ = [("origin_x", C.c_float), ("origin_y", C.c_float), ("origin_z", C.c_float),
                ("orientation_x", C.c_float), ("orientation_y", C.c_float), ("orientation_z", C.c_float),
                ("up_x", C.c_float), ("up_y", C.c_float), ("up_z", C.c_float),
                ("far_plane", C.c_float), ("near_plane", C.c_float), ("fov", C.c_float),
                ("background_start", C.c_float * 3), ("background_end", C.c_float * 3)]


class FlyCam(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("orientation", C.c_float * 3), ("up", C.c_float * 3), ("far_plane", C.c_float),
                ("near_plane", C.c_float), ("fov_degrees", C.c_float), ("speed", C.c_float), ("sensitivity", C.c_float),
                ("yaw", C.c_float), ("pitch", C.c_float)]


KEY_W, KEY_A, KEY_S, KEY_D, KEY_SPACE, KEY_LCTRL, KEY_LSHIFT, KEY_C = 1, 2, 4, 8, 16, 32, 64, 128


class RenderParams(C.Structure):
    _fields_ = [("spp_begin", C.c_uint32), ("spp_count", C.c_uint32), ("spp_stride", C.c_uint32), ("max_depth", C.c_uint32),
                ("seed", C.c_uint64), ("variant", C.c_int32), ("jitter", C.c_int32), ("flags", C.c_uint32),
                ("row_tiles", C.c_uint32)]


class Counters(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("box_tests", C.c_uint64), ("prim_tests", C.c_uint64),
                ("kernel_launches", C.c_uint64)]


class BvhStats(C.Structure):
    _fields_ = [("n_prims", C.c_uint64), ("n_nodes", C.c_uint64), ("max_depth", C.c_uint64), ("smem_nodes", C.c_uint64),
                ("sah_cost", C.c_double), ("build_ms", C.c_double), ("node_bytes", C.c_uint64), ("prim_bytes", C.c_uint64),
                ("h2d_bytes", C.c_uint64), ("builder", C.c_uint64), ("cache_hits", C.c_uint64)]


assert C.sizeof(Hittable) == 32 and C.sizeof(Material) == 16 and C.sizeof(Texture) == 48 and C.sizeof(InputStruct) == 72

# name -> (restype, argtypes); this table is also what tests/test_abi.py checks against include/crt.h
_P = C.c_void_p
SIGNATURES = {
    "crt_context_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "crt_context_destroy": (None, [_P]),
    "crt_last_error_string": (C.c_char_p, [_P]),
    "crt_abi_version": (C.c_int, []),
    "crt_set_stream": (C.c_int, [_P, _P]),
    "crt_synchronize": (C.c_int, [_P]),
    "crt_set_bvh_builder": (C.c_int, [_P, C.c_int32]),
    "crt_set_scene_cache": (C.c_int, [_P, C.c_int32]),
    "crt_scene_upload": (C.c_int, [_P, C.POINTER(SceneDesc)]),
    "crt_scene_from_reference": (C.c_int, [_P, _P]),
    "crt_bvh_get_stats": (C.c_int, [_P, C.POINTER(BvhStats)]),
    "crt_scene_edit_hittable": (C.c_int, [_P, C.c_uint64, C.POINTER(Hittable)]),
    "crt_scene_edit_material": (C.c_int, [_P, C.c_uint64, C.POINTER(Material)]),
    "crt_scene_commit": (C.c_int, [_P, C.c_int, C.POINTER(C.c_int)]),
    "crt_bvh_build_check": (C.c_int, [C.POINTER(SceneDesc), C.POINTER(BvhStats)]),
    "crt_set_camera": (C.c_int, [_P, C.POINTER(InputStruct)]),
    "crt_set_lens": (C.c_int, [_P, C.c_float, C.c_float]),
    "crt_framebuffer_resize": (C.c_int, [_P, C.c_uint32, C.c_uint32]),
    "crt_accum_clear": (C.c_int, [_P]),
    "crt_set_accum_device_buffer": (C.c_int, [_P, _P]),
    "crt_accum_device_ptr": (_P, [_P]),
    "crt_render": (C.c_int, [_P, C.POINTER(RenderParams)]),
    "crt_resolve": (C.c_int, [_P, _P]),
    "crt_last_variant": (C.c_int32, [_P]),
    "crt_resolve_device": (C.c_int, [_P, _P]),
    "crt_resolve_device_truncated": (C.c_int, [_P, _P]),
    "crt_read_accum": (C.c_int, [_P, _P]),
    "crt_read_accum_static": (C.c_int, [_P, _P, C.c_uint32, _P]),
    "crt_resolve_rows": (C.c_int, [_P, C.c_uint32, _P]),
    "crt_row_tiles_rows": (C.c_uint32, [C.c_uint32, C.c_uint32]),
    "crt_read_rgba": (C.c_int, [_P, _P]),
    "crt_rgba_device_ptr": (_P, [_P]),
    "crt_peer_enable": (C.c_int, [_P, C.c_int32]),
    "crt_ipc_export": (C.c_int, [_P, C.c_int32, _P]),
    "crt_ipc_open": (C.c_int, [_P, _P, C.POINTER(_P)]),
    "crt_ipc_close": (C.c_int, [_P, _P]),
    "crt_resolve_peers": (C.c_int, [_P, C.POINTER(_P), C.c_int32, C.c_uint32, C.c_uint32, _P, _P]),
    "crt_primary_ids": (C.c_int, [_P, C.c_int32, C.c_uint64, _P]),
    "crt_get_counters": (C.c_int, [_P, C.POINTER(Counters)]),
    "crt_reset_counters": (C.c_int, [_P]),
    "crt_last_render_ms": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "crt_scene_generate": (C.c_int, [C.c_int32, C.c_uint64, C.c_uint64, C.POINTER(_P)]),
    "crt_scene_builder_desc": (C.POINTER(SceneDesc), [_P]),
    "crt_scene_builder_camera": (C.c_int, [_P, C.c_uint32, C.c_uint32, C.POINTER(InputStruct)]),
    "crt_scene_builder_free": (None, [_P]),
    "crt_camera_look": (C.c_int, [C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_float, C.c_float,
                                  C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(InputStruct)]),
    "crt_flycam_init": (C.c_int, [C.POINTER(FlyCam), C.POINTER(C.c_float)]),
    "crt_flycam_step": (C.c_int, [C.POINTER(FlyCam), C.c_uint32, C.c_float, C.c_float, C.c_float, C.POINTER(C.c_int)]),
    "crt_flycam_inputs": (C.c_int, [C.POINTER(FlyCam), C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(InputStruct)]),
    "crt_load_ppm": (C.c_int, [C.c_char_p, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "crt_decode_jpeg": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                  C.POINTER(C.c_int32)]),
    "crt_load_image": (C.c_int, [C.c_char_p, C.POINTER(C.POINTER(C.c_uint8)), C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                 C.POINTER(C.c_int32)]),
    "crt_free": (None, [_P]),
    "crt_write_ppm": (C.c_int, [C.c_char_p, _P, C.c_uint32, C.c_uint32, C.c_int]),
    "crt_write_png": (C.c_int, [C.c_char_p, _P, C.c_uint32, C.c_uint32, C.c_int]),
    # include/crt_legacy.h
    "LaunchKernel": (None, [_P, C.c_uint, C.c_uint, C.c_uint, C.c_uint, _P, _P, InputStruct]),
    "LaunchRandInit": (None, [_P]),
    "crt_legacy_counters": (C.c_int, [C.POINTER(Counters)]),
    "crt_legacy_shutdown": (None, []),
}

_lib = None
_scenegen = None
SCENEGEN_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libcrt_scenegen.so")
SCENEGEN_SYMBOLS = ("crt_scene_generate", "crt_scene_builder_desc", "crt_scene_builder_camera", "crt_scene_builder_free",
                    "crt_camera_look", "crt_last_error_string", "crt_load_image", "crt_free")


def load_scenegen() -> C.CDLL:
    """The HOST-ONLY library (scene generators + image decoder, no CUDA): for processes that must not map libcrt.so, i.e. the
    CPU arm of bench.py that times the reference's own code."""
    global _scenegen
    if _scenegen is None:
        if not os.path.exists(SCENEGEN_PATH):
            raise RuntimeError(f"{SCENEGEN_PATH} is missing: build it with `make -C cudaraytracer_b200/csrc`")
        lib = C.CDLL(SCENEGEN_PATH)
        for name in SCENEGEN_SYMBOLS:
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = SIGNATURES[name]
        _scenegen = lib
    return _scenegen


def load() -> C.CDLL:
    """Loads libcrt.so and attaches the prototypes. Raises if the library is missing: there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `make -C cudaraytracer_b200/csrc` "
                           "(or __graft_entry__.build()). This package has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
