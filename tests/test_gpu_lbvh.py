"""GPU tests of the device-side BVH builders (csrc/lbvh.cuh: Morton LBVH): the tree is different from the
host SAH tree, the closest hit is not. Primary-hit IDs bit-exact against the oracle, accumulation buffers bit-identical to the host-built tree's, edits
(refit and topology change) on a GPU-built tree, fall-back for degenerate scenes. All through the C ABI."""
import numpy as np
import pytest

import oracle_lib as O
from cudaraytracer_b200 import Renderer, Scene, capi, look_camera

pytestmark = pytest.mark.gpu

SCENES = [("rtiow", capi.SCENE_RTIOW, 0, 160, 90), ("default_rects", capi.SCENE_REFERENCE_DEFAULT, 0, 120, 80),
          ("stress10k", capi.SCENE_STRESS_10K, 0, 128, 72), ("textured", capi.SCENE_TEXTURED, 512, 128, 72)]


def _setup(r, scene, cam, W, H):
    r.upload(scene)
    r.set_camera(cam)
    r.resize(W, H)


GPU_BUILDERS = [capi.BUILDER_GPU_LBVH]


@pytest.mark.parametrize("builder", GPU_BUILDERS, ids=["lbvh"])
@pytest.mark.parametrize("name,kind,hint,W,H", SCENES)
def test_gpu_built_tree_same_hits_as_host_tree(name, kind, hint, W, H, builder):
    s = Scene.generate(kind, 1984, hint)
    cam = s.default_camera(W, H)
    host, gpu = Renderer(0), Renderer(0)
    try:
        host.set_bvh_builder(capi.BUILDER_HOST_SAH)
        gpu.set_bvh_builder(builder)
        _setup(host, s, cam, W, H)
        _setup(gpu, s, cam, W, H)
        sh, sg = host.bvh_stats(), gpu.bvh_stats()
        assert sh.builder == capi.BUILDER_HOST_SAH and sg.builder == builder
        assert sg.n_prims == sh.n_prims and sg.n_nodes == sg.n_prims - 1 and 0 < sg.max_depth < 64
        print(f"{name}: SAH statistic host {sh.sah_cost:.3f} vs GPU builder {builder} {sg.sah_cost:.3f}; depth {sh.max_depth} vs {sg.max_depth}")
        assert sg.sah_cost < 3.0 * sh.sah_cost
        ids = gpu.primary_ids()
        assert np.array_equal(ids, host.primary_ids())
        if s.n_hittables <= 20000:
            ref = O.oracle_render(s, cam, W, H, 0, 5, want_ids=True, want_rgba=False, want_mean=False)["ids"]
            assert np.array_equal(ids, ref)
        for variant in (capi.VARIANT_MEGAKERNEL, capi.VARIANT_PIPELINE, capi.VARIANT_WAVEFRONT):
            host.clear(); gpu.clear()
            host.reset_counters(); gpu.reset_counters()
            host.render(spp=4, max_depth=50, variant=variant)
            gpu.render(spp=4, max_depth=50, variant=variant)
            a, b = host.read_accum(), gpu.read_accum()
            assert host.counters().rays == gpu.counters().rays
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), f"{(a != b).any(axis=-1).sum()} pixels differ ({name}, variant {variant})"
    finally:
        host.close(); gpu.close()


def test_gpu_built_large_grid_ids_vs_brute_force():
    """90 k spheres (a size AUTO gives to the GPU builder): IDs against the oracle's brute-force closest hit on a pixel subset."""
    s = Scene.generate(capi.SCENE_GRID_1M, 1984, 300)
    W, H = 96, 54
    cam = s.default_camera(W, H)
    r = Renderer(0)
    try:
        _setup(r, s, cam, W, H)
        st = r.bvh_stats()
        assert st.n_prims == 300 * 300 + 1 and st.builder == capi.BUILDER_GPU_LBVH  # AUTO: >= 65536 hittables
        ids = r.primary_ids()
        ref = O.oracle_render(s, cam, W, H, 0, 5, want_ids=True, want_rgba=False, want_mean=False, accel=O.BRUTE_FORCE)["ids"]
        assert np.array_equal(ids, ref)
        assert len(np.unique(ids)) > 500
    finally:
        r.close()


@pytest.mark.parametrize("builder", GPU_BUILDERS, ids=["lbvh"])
def test_edits_on_a_gpu_built_tree(builder):
    """Refit (moved / resized spheres, material switch) and a topology change (deactivate, type change) with the GPU builder."""
    s = Scene.generate(capi.SCENE_RTIOW, 1984)
    W, H = 128, 72
    cam = s.default_camera(W, H)
    r, fresh = Renderer(0), Renderer(0)
    try:
        r.set_bvh_builder(builder)
        _setup(r, s, cam, W, H)
        h = s.hittable_array().copy()
        edited = []
        for i in (1, 5, 17, 100, 250, 480):
            f = dict(type=int(h["type"][i]), active=1, cx=float(h["cx"][i]) + 0.3, cy=float(h["cy"][i]) + 0.2, cz=float(h["cz"][i]) - 0.1,
                     a=float(h["a"][i]) * 1.5, b=0.0, material=int(h["material"][(i + 3) % len(h)]))
            r.edit_hittable(i, **f)
            edited.append((i, f))
        assert r.commit() is False  # refit
        f2 = dict(type=capi.SPHERE, active=0, cx=0.0, cy=0.0, cz=0.0, a=1.0, b=0.0, material=int(h["material"][7]))
        r.edit_hittable(7, **f2)
        edited.append((7, f2))
        assert r.commit() is True   # rebuild, on the GPU
        assert r.bvh_stats().builder == builder and r.bvh_stats().n_prims == len(h) - 1
        # the same scene built from scratch by the host builder
        hs = [dict(type=int(x["type"]), active=int(x["active"]), cx=float(x["cx"]), cy=float(x["cy"]), cz=float(x["cz"]), a=float(x["a"]),
                   b=float(x["b"]), material=int(x["material"])) for x in h]
        for i, f in edited:
            hs[i] = f
        d = s.desc
        mats = [dict(type=d.materials[i].type, texture=d.materials[i].texture, param=d.materials[i].param,
                     light_intensity=d.materials[i].light_intensity) for i in range(d.n_materials)]
        texs = [dict(type=d.textures[i].type, image=d.textures[i].image, color=tuple(d.textures[i].color), odd=tuple(d.textures[i].odd),
                     even=tuple(d.textures[i].even)) for i in range(d.n_textures)]
        s2 = Scene.from_arrays(hs, mats, texs)
        fresh.set_bvh_builder(capi.BUILDER_HOST_SAH)
        _setup(fresh, s2, cam, W, H)
        assert np.array_equal(r.primary_ids(), fresh.primary_ids())
        r.clear()
        r.render(spp=4, max_depth=50)
        fresh.render(spp=4, max_depth=50)
        assert np.array_equal(r.read_accum().view(np.uint32), fresh.read_accum().view(np.uint32))
    finally:
        r.close(); fresh.close()


@pytest.mark.parametrize("builder", GPU_BUILDERS, ids=["lbvh"])
def test_degenerate_scenes_fall_back_to_the_host_builder(builder):
    cam = look_camera((0, 0, 5), (0, 0, -1))
    tex = [dict(type=capi.CONSTANT, color=(0.5, 0.5, 0.5))]
    mat = [dict(type=capi.LAMBERTIAN, texture=0)]
    r = Renderer(0)
    try:
        r.set_bvh_builder(builder)
        for hs in ([], [dict(type=capi.SPHERE, cx=0, cy=0, cz=0, a=1.0, material=0, active=1)]):
            sc = Scene.from_arrays(hs, mat, tex)
            _setup(r, sc, cam, 40, 30)
            assert r.bvh_stats().builder == capi.BUILDER_HOST_SAH
            ids = r.primary_ids()
            assert (ids == -1).all() if not hs else (ids.max() == 0 and ids.min() == -1)
        # coincident centres: equal Morton keys are told apart by their sorted position
        same = [dict(type=capi.SPHERE, cx=0, cy=0, cz=0, a=0.2 + 0.1 * k, material=0, active=1) for k in range(9)]
        sc = Scene.from_arrays(same, mat, tex)
        _setup(r, sc, cam, 40, 30)
        assert r.bvh_stats().builder == builder
        ref = O.oracle_render(sc, cam, 40, 30, 0, 5, want_ids=True, want_rgba=False, want_mean=False)["ids"]
        assert np.array_equal(r.primary_ids(), ref)
        bad = Scene.from_arrays([dict(type=capi.SPHERE, cx=0, cy=0, cz=0, a=1.0, material=3)] * 3, mat, tex)
        with pytest.raises(Exception):
            r.upload(bad)
    finally:
        r.close()
