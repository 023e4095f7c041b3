"""Loader for the CHECKERS under oracle/ (test infrastructure). Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs import this."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from cudaraytracer_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")
REFHOST_SO = os.path.join(ROOT, "oracle", "_ref", "libcrt_refhost.so")
REFKERNEL_SO = os.path.join(ROOT, "oracle", "_ref", "libcrt_refkernel.so")

BVH_REFERENCE, BRUTE_FORCE = 0, 1

_oracle = None
_refhost = None


def oracle():
    global _oracle
    if _oracle is None:
        lib = C.CDLL(ORACLE_SO)
        lib.oracle_render.restype = C.c_int
        lib.oracle_render.argtypes = [C.POINTER(capi.SceneDesc), C.POINTER(capi.InputStruct), C.c_uint32, C.c_uint32, C.c_uint32,
                                      C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.POINTER(C.c_uint64)]
        lib.oracle_resolve.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        lib.oracle_curand_uniform.argtypes = [C.c_uint64, C.c_int, C.c_void_p]
        lib.oracle_sphere_hit.restype = C.c_int
        lib.oracle_sphere_hit.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p]
        lib.oracle_aabb_hit.restype = C.c_int
        lib.oracle_aabb_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float]
        lib.oracle_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.oracle_set_pixel_stride.argtypes = [C.c_uint32, C.c_uint32]
        lib.oracle_set_lens.argtypes = [C.c_float, C.c_float]
        lib.oracle_set_seed_base.argtypes = [C.c_uint64]
        _oracle = lib
    return _oracle


def have_refhost() -> bool:
    return os.path.exists(REFHOST_SO)


def refhost():
    global _refhost
    if _refhost is None:
        lib = C.CDLL(REFHOST_SO)
        lib.refhost_render.restype = C.c_int
        lib.refhost_render.argtypes = [C.POINTER(capi.SceneDesc), C.POINTER(capi.InputStruct), C.c_uint32, C.c_uint32, C.c_uint32,
                                       C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
        lib.refhost_curand_uniform.argtypes = [C.c_uint64, C.c_int, C.c_void_p]
        lib.refhost_struct_sizes.argtypes = [C.c_void_p]
        lib.refhost_sphere_hit.restype = C.c_int
        lib.refhost_sphere_hit.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p]
        lib.refhost_aabb_hit.restype = C.c_int
        lib.refhost_aabb_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float]
        _refhost = lib
    return _refhost


def _render(which, scene, cam, width, height, spp, max_depth, jitter_center=False, threads=0, accel=BVH_REFERENCE,
            want_rgba=True, want_mean=True, want_ids=False):
    threads = threads or (os.cpu_count() or 1)
    rgba = np.zeros((height, width), dtype=np.uint32) if want_rgba else None
    mean = np.zeros((height, width, 3), dtype=np.float32) if want_mean else None
    ids = np.full((height, width), -9, dtype=np.int32) if want_ids else None
    rays = C.c_uint64(0)
    p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
    if which == "oracle":
        rc = oracle().oracle_render(C.byref(scene.desc), C.byref(cam), width, height, spp, max_depth, int(jitter_center), threads,
                                    accel, p(rgba), p(mean), p(ids), C.byref(rays))
    else:
        rc = refhost().refhost_render(C.byref(scene.desc), C.byref(cam), width, height, spp, max_depth, int(jitter_center), threads,
                                      p(rgba), p(mean), p(ids), C.byref(rays))
    if rc != 0:
        raise RuntimeError(f"{which}_render failed with {rc}")
    return {"rgba": rgba, "mean": mean, "ids": ids, "rays": int(rays.value)}


def oracle_render(scene, cam, width, height, spp, max_depth, stride=(1, 1), lens=None, **kw):
    """stride = (sx, sy): only pixels with x % sx == 0 and y % sy == 0 (their full-frame rays); the other entries of the
    outputs keep their fill value (ids: -9). lens = (aperture, focus_dist): the thin lens of the new path."""
    lib = oracle()
    lib.oracle_set_pixel_stride(int(stride[0]), int(stride[1]))
    lib.oracle_set_lens(*(lens if lens else (0.0, 1.0)))
    try:
        return _render("oracle", scene, cam, width, height, spp, max_depth, **kw)
    finally:
        lib.oracle_set_pixel_stride(1, 1)
        lib.oracle_set_lens(0.0, 1.0)


def refhost_render(scene, cam, width, height, spp, max_depth, **kw):
    kw.pop("accel", None)
    return _render("refhost", scene, cam, width, height, spp, max_depth, **kw)


def oracle_resolve(accum: np.ndarray) -> np.ndarray:
    accum = np.ascontiguousarray(accum, dtype=np.float32)
    n = accum.size // 4
    out = np.empty(accum.shape[:-1], dtype=np.uint32)
    oracle().oracle_resolve(accum.ctypes.data_as(C.c_void_p), n, out.ctypes.data_as(C.c_void_p))
    return out


# ---- the reference's own CUDA kernel (oracle/_ref/libcrt_refkernel.so; GPU box only) ----
_refkernel = None


def have_refkernel() -> bool:
    return os.path.exists(REFKERNEL_SO)


def refkernel():
    global _refkernel
    if _refkernel is None:
        lib = C.CDLL(REFKERNEL_SO)
        lib.refkernel_render.restype = C.c_int
        lib.refkernel_render.argtypes = [C.POINTER(capi.SceneDesc), C.POINTER(capi.InputStruct), C.c_uint32, C.c_uint32, C.c_uint32,
                                         C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
        lib.refkernel_primary_ids.restype = C.c_int
        lib.refkernel_primary_ids.argtypes = [C.POINTER(capi.SceneDesc), C.POINTER(capi.InputStruct), C.c_uint32, C.c_uint32, C.c_void_p]
        lib.refkernel_primary_ids_ieee.restype = C.c_int
        lib.refkernel_primary_ids_ieee.argtypes = lib.refkernel_primary_ids.argtypes
        _refkernel = lib
    return _refkernel


def refkernel_render(scene, cam, width, height, spp, max_depth, reps=1):
    """Reference CUDA kernel, unmodified (Kernel.cu via LaunchKernel). Returns (rgba8 HxW uint32, [ms per rep])."""
    rgba = np.zeros((height, width), dtype=np.uint32)
    ms = np.zeros(max(reps, 1), dtype=np.float32)
    rc = refkernel().refkernel_render(C.byref(scene.desc), C.byref(cam), width, height, spp, max_depth, reps,
                                      rgba.ctypes.data_as(C.c_void_p), ms.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise RuntimeError(f"refkernel_render failed with {rc}")
    return rgba, [float(x) for x in ms[:reps]]


def refkernel_primary_ids(scene, cam, width, height, ieee=False):
    """The reference's device BVHNode::Hit per pixel; ieee=False: its release flags (-use_fast_math), True: IEEE arithmetic."""
    ids = np.full((height, width), -9, dtype=np.int32)
    fn = refkernel().refkernel_primary_ids_ieee if ieee else refkernel().refkernel_primary_ids
    rc = fn(C.byref(scene.desc), C.byref(cam), width, height, ids.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise RuntimeError(f"refkernel_primary_ids failed with {rc}")
    return ids


def refhost_world(scene):
    """Reference scene graph in host memory -> (handle, root Hittable*) for driving the drop-in LaunchKernel."""
    lib = refhost()
    lib.refhost_world_create.restype = C.c_void_p
    lib.refhost_world_create.argtypes = [C.POINTER(capi.SceneDesc)]
    lib.refhost_world_root.restype = C.c_void_p
    lib.refhost_world_root.argtypes = [C.c_void_p]
    lib.refhost_world_destroy.argtypes = [C.c_void_p]
    h = lib.refhost_world_create(C.byref(scene.desc))
    if not h:
        raise RuntimeError("refhost_world_create failed")
    return h, lib.refhost_world_root(h)


def refhost_world_list_root(handle):
    """The same world behind a HittableList root (Hittable.cuh:532-581)."""
    lib = refhost()
    lib.refhost_world_list_root.restype = C.c_void_p
    lib.refhost_world_list_root.argtypes = [C.c_void_p]
    return lib.refhost_world_list_root(handle)


def refhost_world_destroy(handle):
    refhost().refhost_world_destroy(handle)
