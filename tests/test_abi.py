"""The C-ABI library loads without a GPU and exports every symbol include/*.h declares."""
import ctypes as C
import os
import re

import pytest

from cudaraytracer_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+)?[A-Za-z_][A-Za-z0-9_]*\s*\*?\s+\*?\s*([A-Za-z_][A-Za-z0-9_]*)\s*\(", src, flags=re.M)
    return sorted(set(n for n in names if n.startswith("crt_") or n.startswith("Launch")))


def test_headers_declare_something():
    assert len(_declared("crt.h")) >= 28
    assert set(_declared("crt_legacy.h")) >= {"LaunchKernel", "LaunchRandInit", "LaunchRenderInit"}


@pytest.mark.parametrize("header", ["crt.h", "crt_legacy.h"])
def test_every_declared_symbol_is_exported(header):
    lib = C.CDLL(capi.LIB_PATH)
    missing = [n for n in _declared(header) if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_table_covers_the_header():
    declared = set(_declared("crt.h"))
    assert declared <= set(capi.SIGNATURES), declared - set(capi.SIGNATURES)
    lib = capi.load()
    assert lib.crt_abi_version() == 4


def test_struct_layouts():
    assert C.sizeof(capi.Hittable) == 32 and C.sizeof(capi.Material) == 16 and C.sizeof(capi.Texture) == 48
    assert C.sizeof(capi.InputStruct) == 72 and C.sizeof(capi.RenderParams) == 40 and C.sizeof(capi.Counters) == 40


def test_scene_generators_run_on_host():
    from cudaraytracer_b200 import Scene
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    assert 470 <= s.n_hittables <= 490
    h = s.hittable_array()
    assert h["type"].max() == capi.SPHERE and h["a"][0] == 1000.0
    s2 = Scene.generate(capi.SCENE_RTIOW, 1984)
    assert (s2.hittable_array() == h).all()  # deterministic
    d = Scene.generate(capi.SCENE_REFERENCE_DEFAULT, 1984)
    assert d.n_hittables == 17 and d.hittable_array()["type"][0] == capi.XZRECT
    g = Scene.generate(capi.SCENE_GRID_1M, 1984, 30)
    assert g.n_hittables == 901


def test_no_gpu_fails_loudly():
    """Without a CUDA device context creation must fail with an error, never fall back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = capi.load()
    ctx = C.c_void_p()
    st = lib.crt_context_create(0, C.byref(ctx))
    assert st != capi.CRT_OK and not ctx.value
    assert b"no CUDA device" in lib.crt_last_error_string(None)


def test_c_host_example_builds_and_fails_loudly_without_a_gpu(tmp_path):
    """examples/headless_render.c: a plain-C host over include/crt.h (the reference's host is C++; the boundary is C)."""
    import subprocess
    import torch
    exe = os.path.join(ROOT, "examples", "headless_render")
    assert os.path.exists(exe), "built by make -C cudaraytracer_b200/csrc (target examples)"
    out = str(tmp_path / "f.ppm")
    r = subprocess.run([exe, "1", "64", "36", "2", "8", "2", out], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stderr
        data = open(out, "rb").read()
        assert data.startswith(b"P6\n64 36\n255\n") and len(data) == len(b"P6\n64 36\n255\n") + 64 * 36 * 3
    else:
        assert r.returncode != 0 and "no CUDA device" in r.stderr


def test_native_multigpu_host_example_builds_and_fails_loudly_without_a_gpu(tmp_path):
    """examples/multigpu_render.cpp: one thread and one context per GPU over the C ABI. Without a device every thread must get
    the library's loud error and the process must end (no thread left waiting at a barrier)."""
    import subprocess
    import torch
    exe = os.path.join(ROOT, "examples", "multigpu_render")
    assert os.path.exists(exe), "built by make -C cudaraytracer_b200/csrc (target examples)"
    if torch.cuda.is_available():
        pytest.skip("with a GPU this is tests/test_gpu_multigpu.py::test_native_multigpu_host_example")
    for mode in ("tiles", "spp"):
        r = subprocess.run([exe, "3", mode, "1", "64", "36", "2", "8", str(tmp_path / "f.png")], capture_output=True, text=True, timeout=60)
        assert r.returncode == 1 and "no CUDA device" in r.stderr
    assert subprocess.run([exe, "99"], capture_output=True, text=True, timeout=60).returncode == 2


@pytest.mark.parametrize("kind,hint,n", [(capi.SCENE_RTIOW, 0, None), (capi.SCENE_REFERENCE_DEFAULT, 0, 17), (capi.SCENE_TEXTURED, 64, None),
                                         (capi.SCENE_STRESS_10K, 0, 10001), (capi.SCENE_GRID_1M, 300, 90001)])
def test_host_bvh_build_invariants(kind, hint, n):
    """The host BVH builder (csrc/scene_pack.cpp, multi-threaded above 32k primitives) on every scene kind: every primitive in
    exactly one leaf and inside all ancestor boxes."""
    from cudaraytracer_b200 import Scene
    s = Scene.generate(kind, 1984, hint)
    st = capi.BvhStats()
    rc = capi.load().crt_bvh_build_check(C.byref(s.desc), C.byref(st))
    assert rc == capi.CRT_OK, capi.load().crt_last_error_string(None)
    assert st.n_prims == (n or s.n_hittables) and st.n_nodes == max(st.n_prims - 1, 1)
    assert st.max_depth < 64 and st.sah_cost > 0


def test_host_bvh_build_is_deterministic_and_handles_degenerate_input():
    from cudaraytracer_b200 import Scene
    tex = [dict(type=capi.CONSTANT, color=(0.5, 0.5, 0.5))]
    mat = [dict(type=capi.LAMBERTIAN, texture=0)]
    same = [dict(type=capi.SPHERE, cx=1.0, cy=2.0, cz=3.0, a=0.5, material=0) for _ in range(100)]  # 100 coincident spheres
    s = Scene.from_arrays(same, mat, tex)
    st = capi.BvhStats()
    assert capi.load().crt_bvh_build_check(C.byref(s.desc), C.byref(st)) == capi.CRT_OK
    assert st.n_prims == 100 and st.max_depth <= 9
    for n in (0, 1, 2, 3):
        s = Scene.from_arrays(same[:n], mat, tex)
        assert capi.load().crt_bvh_build_check(C.byref(s.desc), C.byref(st)) == capi.CRT_OK and st.n_prims == n


def test_host_bvh_build_survives_random_and_rejects_malformed_scenes():
    """Randomised scenes (duplicates, collinear centres, 1e-30 and 1e6 sizes, rects, inactive entries) keep the invariants;
    malformed ones (bad indices or types, NaN / infinite geometry) are refused with CRT_ERR_INVALID, never crash or build a
    poisoned tree; an inactive hittable may hold anything."""
    import random
    from cudaraytracer_b200 import Scene
    lib = capi.load()
    tex = [dict(type=capi.CONSTANT, color=(0.5, 0.5, 0.5))]
    mat = [dict(type=capi.LAMBERTIAN, texture=0), dict(type=capi.METAL, texture=0), dict(type=capi.DIELECTRIC, texture=0)]
    rng = random.Random(7)

    def val(scale):
        k = rng.random()
        return 0.0 if k < 0.05 else scale * 1e-30 if k < 0.10 else scale * 1e6 if k < 0.15 else rng.uniform(-scale, scale)

    for it in range(120):
        n = rng.choice([0, 1, 2, 3, 5, 8, 9, 17, 33, 64, 257])
        mode = rng.choice(["uniform", "line", "dup", "mixed", "huge"])
        hs = []
        for i in range(n):
            t = rng.choice([capi.SPHERE] * 4 + [capi.XYRECT, capi.XZRECT, capi.YZRECT])
            c = {"line": (float(i), 0.0, 0.0), "dup": (float(i % 3), 1.0, 2.0), "huge": (val(1e5), val(1e5), val(1e5))}.get(
                mode, (rng.uniform(-10, 10), rng.uniform(-10, 10), rng.uniform(-10, 10)))
            a = abs(val(2.0)) if mode in ("mixed", "huge") else rng.uniform(0.01, 1.0)
            hs.append(dict(type=t, cx=c[0], cy=c[1], cz=c[2], a=-a if rng.random() < 0.1 else a, b=abs(val(3.0)),
                           material=rng.randrange(3), active=int(rng.random() > 0.1)))
        st = capi.BvhStats()
        s = Scene.from_arrays(hs, mat, tex)
        assert lib.crt_bvh_build_check(C.byref(s.desc), C.byref(st)) == capi.CRT_OK, (it, lib.crt_last_error_string(None))
        assert st.n_prims == sum(h["active"] for h in hs) and st.max_depth < 64

    base = dict(type=capi.SPHERE, cx=0.0, cy=0.0, cz=0.0, a=1.0, b=0.0, material=0, active=1)
    others = [dict(base, cx=float(i + 3)) for i in range(5)]
    nan, inf = float("nan"), float("inf")
    bad = [dict(base, cx=nan), dict(base, cy=inf), dict(base, cz=-inf), dict(base, a=nan), dict(base, a=inf),
           dict(base, type=capi.XZRECT, b=nan), dict(base, material=5), dict(base, material=-1), dict(base, type=17)]
    for h in bad:
        for rest in ([], others):
            s = Scene.from_arrays([h] + rest, mat, tex)
            assert lib.crt_bvh_build_check(C.byref(s.desc), C.byref(capi.BvhStats())) == capi.CRT_ERR_INVALID, h
            assert b"hittable 0" in lib.crt_last_error_string(None)
    st = capi.BvhStats()
    s = Scene.from_arrays([dict(base, cx=nan, a=inf, material=99, active=0)] + others, mat, tex)  # inactive: not looked at
    assert lib.crt_bvh_build_check(C.byref(s.desc), C.byref(st)) == capi.CRT_OK and st.n_prims == 5
    s = Scene.from_arrays([dict(base, a=-0.5)], mat, tex)  # negative radius = the hollow-sphere idiom: its box uses |r|
    assert lib.crt_bvh_build_check(C.byref(s.desc), C.byref(st)) == capi.CRT_OK and st.n_prims == 1
    for m, t in (([dict(type=capi.LAMBERTIAN, texture=3)], tex), ([dict(type=9, texture=0)], tex),
                 ([dict(type=capi.LAMBERTIAN, texture=0)], [dict(type=capi.IMAGE, image=2)])):
        s = Scene.from_arrays([base], m, t)
        assert lib.crt_bvh_build_check(C.byref(s.desc), C.byref(capi.BvhStats())) == capi.CRT_ERR_INVALID

