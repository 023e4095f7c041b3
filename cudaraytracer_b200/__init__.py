"""B200-native replacement of CudaRayTracer's render path.

The product is ``cudaraytracer_b200/libcrt.so`` (C ABI in include/crt.h, CUDA kernels for sm_100a).
This package only loads it through ctypes; it has no CPU fallback and imports nothing from oracle/.
"""
from . import capi
from .renderer import CrtError, Renderer, Scene, load_image, look_camera, unpack_rgba8, write_png, write_ppm

__all__ = ["capi", "CrtError", "Renderer", "Scene", "load_image", "look_camera", "unpack_rgba8", "write_png", "write_ppm"]
