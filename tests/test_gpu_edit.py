"""GPU tests of the scene-edit path: staged edits + crt_scene_commit = GPU refit (topology kept) or host
rebuild (topology changed). Bar: after a commit the renderer is indistinguishable from a fresh upload of the edited scene —
primary-hit IDs bit-exact vs the oracle, accumulation buffers bit-identical to a fresh context."""
import ctypes as C
import time

import numpy as np
import pytest

import oracle_lib as O
from cudaraytracer_b200 import Renderer, Scene, capi

pytestmark = pytest.mark.gpu


def _lists(scene):
    d = scene.desc
    hs = [dict(type=h.type, active=h.active, cx=h.cx, cy=h.cy, cz=h.cz, a=h.a, b=h.b, material=h.material)
          for h in (d.hittables[i] for i in range(d.n_hittables))]
    ms = [dict(type=m.type, texture=m.texture, param=m.param, light_intensity=m.light_intensity) for m in (d.materials[i] for i in range(d.n_materials))]
    ts = [dict(type=t.type, image=t.image, color=tuple(t.color), odd=tuple(t.odd), even=tuple(t.even)) for t in (d.textures[i] for i in range(d.n_textures))]
    return hs, ms, ts


def _fresh(scene, cam, W, H, spp=4):
    r = Renderer(0)
    r.upload(scene); r.set_camera(cam); r.resize(W, H)
    ids = r.primary_ids()
    r.render(spp=spp, max_depth=50)
    acc = r.read_accum()
    r.close()
    return ids, acc


def test_refit_and_rebuild_match_a_fresh_upload():
    W, H = 200, 113
    base = Scene.generate(capi.SCENE_RTIOW, 1984)
    cam = base.default_camera(W, H)
    hs, ms, ts = _lists(base)
    r = Renderer(0)
    r.upload(base); r.set_camera(cam); r.resize(W, H)
    rng = np.random.default_rng(11)
    # 1. move / resize 80 spheres, re-assign some materials: topology kept -> refit
    for i in rng.choice(np.arange(1, len(hs)), 80, replace=False):
        h = hs[int(i)]
        h["cx"] += float(rng.uniform(-0.6, 0.6)); h["cz"] += float(rng.uniform(-0.6, 0.6)); h["cy"] += float(rng.uniform(0, 0.4))
        h["a"] *= float(rng.uniform(0.6, 1.6))
        if rng.random() < 0.3: h["material"] = int(rng.integers(0, len(ms)))
        r.edit_hittable(int(i), **h)
    assert r.commit() is False
    edited = Scene.from_arrays(hs, ms, ts)
    ref_ids = O.oracle_render(edited, cam, W, H, 0, 50, want_ids=True, want_rgba=False, want_mean=False)["ids"]
    ids = r.primary_ids()
    assert np.array_equal(ids, ref_ids), f"{(ids != ref_ids).sum()} IDs differ from the oracle after the refit"
    r.render(spp=4, max_depth=50)
    f_ids, f_acc = _fresh(edited, cam, W, H)
    assert np.array_equal(f_ids, ids)
    assert np.array_equal(r.read_accum().view(np.uint32), f_acc.view(np.uint32)), "refit scene renders differently from a fresh upload"
    # 2. edit a MATERIAL (type change Lambertian -> Metal): every primitive that uses it gets a new record and leaf code
    ms[hs[5]["material"]] = dict(type=capi.METAL, texture=ms[hs[5]["material"]]["texture"], param=0.1, light_intensity=0)
    r.edit_material(hs[5]["material"], **ms[hs[5]["material"]])
    assert r.commit() is False
    r.clear(); r.render(spp=4, max_depth=50)
    edited = Scene.from_arrays(hs, ms, ts)
    _, f_acc = _fresh(edited, cam, W, H)
    assert np.array_equal(r.read_accum().view(np.uint32), f_acc.view(np.uint32))
    # 3. delete three spheres (isActive = false, CudaLayer.cpp:1372-1387) and turn one into a rect: topology changes -> rebuild
    for i in (7, 100, 300):
        hs[i]["active"] = 0
        r.edit_hittable(i, **hs[i])
    hs[50].update(type=capi.XZRECT, a=1.5, b=1.5, cy=0.7)
    r.edit_hittable(50, **hs[50])
    assert r.commit() is True
    edited = Scene.from_arrays(hs, ms, ts)
    ref_ids = O.oracle_render(edited, cam, W, H, 0, 50, want_ids=True, want_rgba=False, want_mean=False)["ids"]
    ids = r.primary_ids()
    assert np.array_equal(ids, ref_ids) and not np.isin(ids, [7, 100, 300]).any()
    r.clear(); r.render(spp=4, max_depth=50)
    _, f_acc = _fresh(edited, cam, W, H)
    assert np.array_equal(r.read_accum().view(np.uint32), f_acc.view(np.uint32))
    # 4. nothing staged: commit is a no-op; bad edits are refused
    assert r.commit() is False
    from cudaraytracer_b200 import CrtError
    with pytest.raises(CrtError):
        r.edit_hittable(len(hs), **hs[0])
    with pytest.raises(CrtError):
        r.edit_hittable(0, **dict(hs[0], material=len(ms)))
    r.close()


def test_refit_of_a_large_scene_is_much_cheaper_than_a_rebuild():
    """300 x 300 grid (90 001 spheres): moving 2000 of them and refitting on the GPU vs rebuilding on the host, and vs
    rebuilding on the GPU (lbvh.cuh), which must beat the host rebuild as well."""
    W, H = 160, 90
    s = Scene.generate(capi.SCENE_GRID_1M, 1984, 300)
    cam = s.default_camera(W, H)
    hs, ms, ts = _lists(s)
    r = Renderer(0)
    r.set_bvh_builder(capi.BUILDER_HOST_SAH)
    r.upload(s); r.set_camera(cam); r.resize(W, H)
    r.edit_hittable(1, **hs[1])
    assert r.commit() is False  # first use of the refit kernels (lazy module load) stays out of the timing
    rng = np.random.default_rng(2)
    t_refit = 1e9
    for rep in range(3):  # best of three: the box's host cores are shared
        for i in rng.choice(np.arange(1, len(hs)), 2000, replace=False):
            h = hs[int(i)]
            h["cy"] += float(rng.uniform(0.0, 1.0)); h["a"] *= float(rng.uniform(0.8, 1.5))
            r.edit_hittable(int(i), **h)
        t0 = time.perf_counter(); rebuilt = r.commit(); t_refit = min(t_refit, time.perf_counter() - t0)
    assert rebuilt is False
    ids = r.primary_ids()
    edited = Scene.from_arrays(hs, ms, ts)
    ref_ids = O.oracle_render(edited, cam, W, H, 0, 50, want_ids=True, want_rgba=False, want_mean=False, accel=O.BRUTE_FORCE)["ids"]
    assert np.array_equal(ids, ref_ids)
    r.commit(force_rebuild=True)
    t0 = time.perf_counter(); assert r.commit(force_rebuild=True) is True; t_rebuild = time.perf_counter() - t0
    assert np.array_equal(r.primary_ids(), ref_ids)
    r.set_bvh_builder(capi.BUILDER_GPU_LBVH)
    r.commit(force_rebuild=True)  # first use: allocations, cub's kernels
    t0 = time.perf_counter(); assert r.commit(force_rebuild=True) is True; t_gpu = time.perf_counter() - t0
    assert r.bvh_stats().builder == capi.BUILDER_GPU_LBVH
    assert np.array_equal(r.primary_ids(), ref_ids)
    print(f"refit {t_refit * 1e3:.2f} ms vs host rebuild {t_rebuild * 1e3:.2f} ms vs GPU rebuild {t_gpu * 1e3:.2f} ms")
    assert t_refit < t_rebuild and t_gpu < t_rebuild
    r.close()


def test_unchanged_scene_upload_is_a_cache_hit_and_any_change_is_not():
    """crt_scene_upload hashes the description: the scene already on the device is not validated, built or copied again
    (crt_bvh_stats.cache_hits / h2d_bytes say so); a changed sphere, an edit + commit, or the switch turned off all upload."""
    W, H = 160, 90
    base = Scene.generate(capi.SCENE_RTIOW, 1984)
    cam = base.default_camera(W, H)
    hs, ms, ts = _lists(base)
    r = Renderer(0)
    try:
        r.upload(base); r.set_camera(cam); r.resize(W, H)
        st = r.bvh_stats()
        assert st.h2d_bytes > 0 and st.cache_hits == 0
        r.render(spp=3, max_depth=50)
        first = r.read_accum().view(np.uint32).copy()
        r.upload(base)
        st = r.bvh_stats()
        assert st.h2d_bytes == 0 and st.cache_hits == 1
        same = Scene.from_arrays(hs, ms, ts)  # another object, other addresses, same content
        r.upload(same)
        assert r.bvh_stats().cache_hits == 2
        r.clear(); r.render(spp=3, max_depth=50)
        assert np.array_equal(first, r.read_accum().view(np.uint32))
        hs2 = [dict(h) for h in hs]
        hs2[7]["cy"] += 0.25
        moved = Scene.from_arrays(hs2, ms, ts)
        r.upload(moved)
        st = r.bvh_stats()
        assert st.h2d_bytes > 0 and st.cache_hits == 0
        ref_ids = O.oracle_render(moved, cam, W, H, 0, 50, want_ids=True, want_rgba=False, want_mean=False)["ids"]
        assert np.array_equal(r.primary_ids(), ref_ids)
        # an edit makes the device scene differ from what was hashed: the ORIGINAL description must upload again
        r.upload(base)
        assert r.bvh_stats().h2d_bytes > 0
        r.edit_hittable(7, **hs2[7])
        r.commit()
        r.upload(base)
        assert r.bvh_stats().h2d_bytes > 0 and r.bvh_stats().cache_hits == 0
        r.clear(); r.render(spp=3, max_depth=50)
        assert np.array_equal(first, r.read_accum().view(np.uint32))
        r.set_scene_cache(False)
        r.upload(base)
        assert r.bvh_stats().h2d_bytes > 0 and r.bvh_stats().cache_hits == 0
    finally:
        r.close()


def test_refit_path_validates_the_material_it_packs():
    """a hittable re-pointed (no topology change) at a material that no active hittable used at upload, whose texture
    chain is broken, must be refused by crt_scene_commit instead of being packed (out-of-bounds reads on host and device)."""
    hs = [dict(type=capi.SPHERE, active=1, cx=0.0, cy=-100.5, cz=-1.0, a=100.0, b=0.0, material=0),
          dict(type=capi.SPHERE, active=1, cx=0.0, cy=0.0, cz=-1.0, a=0.5, b=0.0, material=0),
          dict(type=capi.SPHERE, active=1, cx=1.0, cy=0.0, cz=-1.0, a=0.5, b=0.0, material=0)]
    ms = [dict(type=capi.LAMBERTIAN, texture=0), dict(type=capi.LAMBERTIAN, texture=7),   # texture index out of range, unused
          dict(type=capi.LAMBERTIAN, texture=1)]                                         # image index out of range, unused
    ts = [dict(type=capi.CONSTANT, color=(0.5, 0.5, 0.5)), dict(type=capi.IMAGE, image=3)]
    s = Scene.from_arrays(hs, ms, ts)
    r = Renderer(0)
    try:
        r.upload(s)
        for bad in (1, 2):
            h = dict(hs[1]); h["material"] = bad
            r.edit_hittable(1, **h)
            with pytest.raises(Exception) as e:
                r.commit()
            assert "material" in str(e.value) or "texture" in str(e.value)
        h = dict(hs[1]); h["cx"] = 0.25
        r.edit_hittable(1, **h)  # a valid edit replaces the pending bad one
        assert r.commit() is False
    finally:
        r.close()
