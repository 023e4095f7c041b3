"""crt_read_accum_static (include/crt.h): the allocation-free variant of crt_read_accum, whose staging array is a static __device__
array inside libcrt.so instead of a cudaMalloc'ed buffer. No counterpart in the reference.
Through the C ABI, against the existing entry point on the same accumulation buffer: the float4 views must be
BIT-IDENTICAL (the same conversion of the same integers; this is the bar of the other accumulation-buffer tests), on two calls
in a row with different buffer contents - the array is shared by all calls and keeps the last call's contents.

This file is named test_zz_* so that pytest collects it LAST: a failure here must not hide
another test. It passes alone in a fresh process and after the rest of the suite (it makes its own context and scene)."""
import ctypes as C

import numpy as np
import pytest

from cudaraytracer_b200 import Renderer, Scene, capi
from cudaraytracer_b200.renderer import CrtError

pytestmark = pytest.mark.gpu


def test_read_accum_static_matches_read_accum_on_two_calls_in_a_row():
    import torch
    W, H = 400, 225  # C1 of BASELINE.json: 90 000 pixels <= CRT_STATIC_ACCUM_MAX_PIXELS
    assert W * H <= capi.STATIC_ACCUM_MAX_PIXELS
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    r = Renderer(0)
    try:
        r.upload(s); r.set_camera(s.default_camera(W, H)); r.resize(W, H)
        r.render(spp=2, max_depth=50)
        r.read_accum_static()  # first use: the module's static array gets its device address (cudaGetSymbolAddress, once per device)
        r.synchronize()
        free_before = torch.cuda.mem_get_info(0)[0]

        # call 1: 2 spp in the buffer
        want1 = r.read_accum()
        got1 = r.read_accum_static().reshape(H, W, 4)
        assert np.array_equal(got1.view(np.uint32), want1.view(np.uint32)), f"{(got1 != want1).any(axis=-1).sum()} pixels differ (call 1)"
        assert (got1[..., 3] == 2).all() and got1[..., :3].max() > 0

        # call 2, right after: other contents (5 spp now) and FEWER pixels, from an offset inside the buffer - the tail of the
        # static array still holds call 1's values and must not show up
        r.render(spp=3, max_depth=50, spp_begin=2)
        want2 = r.read_accum()
        n2, first = 50 * W, 100 * W  # rows 100..149
        got2 = r.read_accum_static(r.accum_ptr() + first * 32, n2).reshape(50, W, 4)
        assert np.array_equal(got2.view(np.uint32), want2[100:150].view(np.uint32)), "call 2 differs from crt_read_accum"
        assert (got2[..., 3] == 5).all()
        assert not np.array_equal(got2, got1[100:150])  # the buffer really changed between the calls

        # and the whole frame again: two full-size calls in a row agree with the existing entry point both times
        got3 = r.read_accum_static().reshape(H, W, 4)
        assert np.array_equal(got3.view(np.uint32), want2.view(np.uint32))

        r.synchronize()
        free_after = torch.cuda.mem_get_info(0)[0]
        assert free_after >= free_before, f"device memory was allocated during the calls: {free_before - free_after} bytes"
    finally:
        r.close()


def test_read_accum_static_rejects_sizes_beyond_the_static_array():
    W, H = 64, 36
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    r = Renderer(0)
    try:
        r.upload(s); r.set_camera(s.default_camera(W, H)); r.resize(W, H)
        out = np.zeros(4, dtype=np.float32)
        lib = capi.load()
        st = lib.crt_read_accum_static(r.ctx, C.c_void_p(r.accum_ptr()), capi.STATIC_ACCUM_MAX_PIXELS + 1, out.ctypes.data_as(C.c_void_p))
        assert st == capi.CRT_ERR_INVALID
        msg = lib.crt_last_error_string(r.ctx).decode()
        assert "CRT_STATIC_ACCUM_MAX_PIXELS" in msg and "262144" in msg
        with pytest.raises(CrtError, match="CRT_STATIC_ACCUM_MAX_PIXELS"):  # the same refusal through the Python host
            r.read_accum_static(r.accum_ptr(), capi.STATIC_ACCUM_MAX_PIXELS + 1)
        assert lib.crt_read_accum_static(r.ctx, None, 4, out.ctypes.data_as(C.c_void_p)) == capi.CRT_ERR_INVALID   # NULL input
        assert lib.crt_read_accum_static(r.ctx, C.c_void_p(r.accum_ptr()), 4, None) == capi.CRT_ERR_INVALID        # NULL output
        assert lib.crt_read_accum_static(r.ctx, C.c_void_p(r.accum_ptr() + 8), 1, out.ctypes.data_as(C.c_void_p)) == capi.CRT_ERR_INVALID  # misaligned
        assert lib.crt_read_accum_static(r.ctx, C.c_void_p(r.accum_ptr()), 0, out.ctypes.data_as(C.c_void_p)) == capi.CRT_OK  # nothing to do
        # exactly the maximum is accepted (a 512 x 512 frame)
        r.resize(512, 512)
        r.set_camera(s.default_camera(512, 512))
        r.render(spp=1, max_depth=8)
        assert np.array_equal(r.read_accum_static().reshape(512, 512, 4).view(np.uint32), r.read_accum().view(np.uint32))
    finally:
        r.close()
