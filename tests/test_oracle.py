"""CPU tests of the oracle (oracle/oracle.c): pinned against the reference's own headers compiled for the
host (oracle/_ref/libcrt_refhost.so, present where /root/reference is mounted or the prebuilt .so travelled)
and against the committed golden vectors made from that build (tests/golden/make_golden.py)."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as O
from cudaraytracer_b200 import Scene, capi

GOLD = os.path.join(os.path.dirname(__file__), "golden")
needs_refhost = pytest.mark.skipif(not O.have_refhost(), reason="oracle/_ref/libcrt_refhost.so not built")

SCENES = [("rtiow", capi.SCENE_RTIOW, 0), ("refdefault", capi.SCENE_REFERENCE_DEFAULT, 0), ("textured", capi.SCENE_TEXTURED, 256),
          ("stress", capi.SCENE_STRESS_10K, 20), ("grid", capi.SCENE_GRID_1M, 24)]


def test_xorwow_known_answer():
    # SURVEY.md §8c probe of the toolkit's host path: curand_init(1989,0,0) -> 0.231103912, 0.408436388
    out = np.zeros(2, dtype=np.float32)
    O.oracle().oracle_curand_uniform(1989, 2, out.ctypes.data_as(C.c_void_p))
    assert out[0] == np.float32(0.231103912) and out[1] == np.float32(0.408436388)


def test_philox_known_answer():
    # Random123 kat_vectors for philox4x32-10
    kats = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
            ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
            ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kats:
        c = np.array(ctr, dtype=np.uint32); k = np.array(key, dtype=np.uint32); o = np.zeros(4, dtype=np.uint32)
        O.oracle().oracle_philox4x32_10(c.ctypes.data_as(C.c_void_p), k.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p))
        assert tuple(int(x) for x in o) == want


def test_sphere_unit_case():
    # unit sphere at the origin, ray from (0,0,1.5)... use the documented t=0.5,u=0.25,v=0.5 case
    out = np.zeros(9, dtype=np.float32)
    f = lambda *v: np.array(v, dtype=np.float32)
    c, ro, rd = f(0, 0, 0), f(0, 0, 1.5), f(0, 0, -1)
    hit = O.oracle().oracle_sphere_hit(c.ctypes.data_as(C.c_void_p), 1.0, ro.ctypes.data_as(C.c_void_p), rd.ctypes.data_as(C.c_void_p),
                                       0.001, 3.4e38, out.ctypes.data_as(C.c_void_p))
    assert hit == 1 and out[0] == 0.5 and out[7] == 0.25 and out[8] == 0.5


@needs_refhost
def test_struct_sizes_of_reference():
    sizes = np.zeros(16, dtype=np.int32)
    O.refhost().refhost_struct_sizes(sizes.ctypes.data_as(C.c_void_p))
    # Hittable Sphere BVHNode Material Lambertian Metal Dielectric DiffuseLight Texture Constant Checker Image HitRecord curandState XYRect InputStruct
    assert list(sizes) == [16, 24, 48, 16, 8, 16, 4, 16, 16, 16, 16, 32, 48, 48, 32, 72]


@needs_refhost
def test_xorwow_matches_toolkit_header():
    a = np.zeros(4096, dtype=np.float32); b = np.zeros(4096, dtype=np.float32)
    for seed in (0, 1984, 1984 + 123456, 2**40 + 7):
        O.oracle().oracle_curand_uniform(seed, a.size, a.ctypes.data_as(C.c_void_p))
        O.refhost().refhost_curand_uniform(seed, b.size, b.ctypes.data_as(C.c_void_p))
        assert np.array_equal(a, b)


@needs_refhost
def test_sphere_and_aabb_match_reference_functions():
    rng = np.random.default_rng(7)
    n_hit = 0
    for _ in range(4000):
        c = rng.uniform(-2, 2, 3).astype(np.float32); r = np.float32(rng.uniform(0.1, 1.5))
        ro = rng.uniform(-4, 4, 3).astype(np.float32); rd = rng.uniform(-1, 1, 3).astype(np.float32)
        oa = np.zeros(9, dtype=np.float32); ob = np.zeros(9, dtype=np.float32)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        ha = O.oracle().oracle_sphere_hit(p(c), r, p(ro), p(rd), 0.001, 3.4e38, p(oa))
        hb = O.refhost().refhost_sphere_hit(p(c), r, p(ro), p(rd), 0.001, 3.4e38, p(ob))
        assert ha == hb and np.array_equal(oa, ob)
        n_hit += ha
        lo = c - r; hi = c + r
        assert O.oracle().oracle_aabb_hit(p(lo), p(hi), p(ro), p(rd), 0.001, 3.4e38) == \
            O.refhost().refhost_aabb_hit(p(lo), p(hi), p(ro), p(rd), 0.001, 3.4e38)
    assert n_hit > 100


@needs_refhost
@pytest.mark.parametrize("name,kind,hint", SCENES)
def test_oracle_bit_exact_vs_reference_host_build(name, kind, hint):
    s = Scene.generate(kind, 1984, hint)
    W, H = 96, 54
    cam = s.default_camera(W, H)
    a = O.oracle_render(s, cam, W, H, 4, 50, want_ids=True)
    b = O.refhost_render(s, cam, W, H, 4, 50, want_ids=True)
    assert np.array_equal(a["ids"], b["ids"])
    assert np.array_equal(a["mean"].view(np.uint32), b["mean"].view(np.uint32))
    assert np.array_equal(a["rgba"], b["rgba"])
    assert a["rays"] == b["rays"]


@pytest.mark.parametrize("name,kind,hint", SCENES)
def test_oracle_vs_golden_fixtures(name, kind, hint):
    """Goldens were produced by the reference host build (tests/golden/make_golden.py)."""
    path = os.path.join(GOLD, f"{name}_64x36_spp2.npz")
    g = np.load(path)
    s = Scene.generate(kind, 1984, hint)
    W, H = 64, 36
    cam = s.default_camera(W, H)
    a = O.oracle_render(s, cam, W, H, 2, 50, want_ids=True)
    assert np.array_equal(a["ids"], g["ids"])
    assert np.array_equal(a["rgba"], g["rgba"])
    assert np.array_equal(a["mean"].view(np.uint32), g["mean_bits"])
    assert a["rays"] == int(g["rays"])


def test_brute_force_equals_reference_bvh():
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    W, H = 64, 36
    cam = s.default_camera(W, H)
    a = O.oracle_render(s, cam, W, H, 2, 50, want_ids=True, accel=O.BVH_REFERENCE)
    b = O.oracle_render(s, cam, W, H, 2, 50, want_ids=True, accel=O.BRUTE_FORCE)
    assert np.array_equal(a["ids"], b["ids"]) and np.array_equal(a["rgba"], b["rgba"])


def test_empty_and_inactive_scenes():
    W, H = 16, 9
    empty = Scene.from_arrays([], [], [])
    from cudaraytracer_b200 import look_camera
    cam = look_camera((0, 0, 5), (0, 0, -1))
    a = O.oracle_render(empty, cam, W, H, 1, 5, want_ids=True)
    assert (a["ids"] == -1).all() and a["rays"] == W * H
    tex = [dict(type=capi.CONSTANT, color=(0.5, 0.5, 0.5))]
    mat = [dict(type=capi.LAMBERTIAN, texture=0)]
    one = Scene.from_arrays([dict(type=capi.SPHERE, cx=0, cy=0, cz=0, a=1.0, material=0, active=0)], mat, tex)
    b = O.oracle_render(one, cam, W, H, 1, 5, want_ids=True)
    assert (b["ids"] == -1).all()
    one = Scene.from_arrays([dict(type=capi.SPHERE, cx=0, cy=0, cz=0, a=1.0, material=0, active=1)], mat, tex)
    c = O.oracle_render(one, cam, W, H, 1, 5, want_ids=True)
    assert (c["ids"] == 0).any() and (c["ids"] == -1).any()


def test_pixel_stride_computes_the_full_frame_rays_of_a_subset():
    """oracle_set_pixel_stride (bench.py `parity`, full-size GPU tests): the strided pixels get exactly their full-frame IDs and
    colours, everything else is left alone."""
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    W, H = 96, 54
    cam = s.default_camera(W, H)
    full = O.oracle_render(s, cam, W, H, 2, 50, want_ids=True)
    sub = O.oracle_render(s, cam, W, H, 2, 50, want_ids=True, stride=(8, 6))
    assert np.array_equal(sub["ids"][::6, ::8], full["ids"][::6, ::8])
    assert np.array_equal(sub["rgba"][::6, ::8], full["rgba"][::6, ::8])
    mask = np.ones((H, W), bool)
    mask[::6, ::8] = False
    assert (sub["ids"][mask] == -9).all() and (sub["rgba"][mask] == 0).all()
    again = O.oracle_render(s, cam, W, H, 2, 50, want_ids=True)  # the stride does not stick
    assert np.array_equal(again["ids"], full["ids"])


def test_lens_extension_of_the_oracle():
    """aperture 0 is the reference camera bit for bit (so the pinned oracle stays pinned); an open lens blurs what is off the
    focus plane and leaves the mean radiance alone."""
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    W, H = 64, 36
    cam = s.default_camera(W, H)
    pin = O.oracle_render(s, cam, W, H, 8, 50)
    zero = O.oracle_render(s, cam, W, H, 8, 50, lens=(0.0, 10.0))
    assert np.array_equal(pin["rgba"], zero["rgba"]) and pin["rays"] == zero["rays"]
    wide = O.oracle_render(s, cam, W, H, 8, 50, lens=(0.6, 10.0))
    assert not np.array_equal(pin["rgba"], wide["rgba"])
    assert abs(float(wide["mean"].mean()) - float(pin["mean"].mean())) < 0.05 * float(pin["mean"].mean())
