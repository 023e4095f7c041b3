"""The BVH traversal of the kernels (cudaraytracer_b200/csrc/trace.cuh) compiled for the HOST (tests/sim/trace_sim.cpp, g++ with
-DCRT_HOST_SIM) and run against a brute-force closest hit on the trees the library's own builder packs. No GPU: this checks the
LOGIC of every traversal form - the megakernel's node_step loop on global and on word-major shared tables, the pipelined
kernel's node, node, node, leaf rounds with node_step_all (staged or 'global' nodes) and with node_visit_pair (child-interleaved
staged nodes, FFMA2 emulated with fmaf, sentinel stack, generic and sphere-only leaf visits) - bit for bit; the device arithmetic itself is what the -m gpu parity tests pin.
The pipelined forms must also VISIT the same inner nodes per ray (all fast paths cull with the binary16-rounded boxes: with the
binary32 boxes in one of them, 6 of 3000 rays of the 10 k scene walk different nodes), which is what keeps the variants' frames
bit-identical although the reference's sphere formula reports grazing hits outside a leaf box."""
import os
import subprocess

import pytest

from cudaraytracer_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def trace_sim(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("sim") / "trace_sim")
    libdir = os.path.dirname(capi.LIB_PATH)
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = ["g++", "-std=c++17", "-O1", "-ffp-contract=off", "-Wno-attributes", "-Wno-unknown-pragmas",
           "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "cudaraytracer_b200", "csrc"), "-I", os.path.join(cuda, "include"),
           os.path.join(ROOT, "tests", "sim", "trace_sim.cpp"), "-o", exe, "-L", libdir, "-lcrt", f"-Wl,-rpath,{libdir}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return exe


@pytest.mark.parametrize("kind,hint,rays", [(capi.SCENE_RTIOW, 0, 6000), (capi.SCENE_REFERENCE_DEFAULT, 0, 6000), (capi.SCENE_TEXTURED, 0, 4000),
                                           (capi.SCENE_STRESS_10K, 0, 3000), (capi.SCENE_GRID_1M, 40, 3000), (capi.SCENE_GRID_1M, 1, 2000),
                                           (capi.SCENE_GRID_1M, 2, 2000)])
def test_every_traversal_form_finds_the_brute_force_closest_hit(trace_sim, kind, hint, rays):
    r = subprocess.run([trace_sim, str(kind), str(hint), str(rays), str(kind * 100 + hint)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert " 0 mismatches" in r.stdout


@pytest.mark.parametrize("kind,hint", [(capi.SCENE_RTIOW, 0), (capi.SCENE_TEXTURED, 0), (capi.SCENE_STRESS_10K, 0), (capi.SCENE_GRID_1M, 3)])
def test_rotated_trees_are_still_trees(trace_sim, kind, hint):
    """CRT_BVH_ROTATE (experimental post-pass of the host builder: tree rotations) must leave a tree every traversal form
    agrees on with brute force; tests/sim/path_sim measures what it does to the box / primitive tests per ray."""
    env = dict(os.environ, CRT_BVH_ROTATE="4")
    r = subprocess.run([trace_sim, str(kind), str(hint), "3000", "11"], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and " 0 mismatches" in r.stdout, r.stdout + r.stderr



def test_binary16_helpers_round_half_extents_up_and_decode_exactly(trace_sim):
    """crt_f2h_ru must give the SMALLEST binary16 >= |x| (+inf beyond 65504; a box may grow, never shrink) and crt_h2_to_f2 /
    crt_h16_up must decode it exactly: checked against numpy's float16 on special values and 20 000 random bit patterns. (The
    device forms are cvt.rp.f16.f32 / cvt.f32.f16; the GPU parity tests pin those through the frames.)"""
    import numpy as np
    r = subprocess.run([trace_sim, "half", "20000"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    rows = np.array([[int(w, 16) for w in line.split()] for line in r.stdout.splitlines()], dtype=np.uint64)
    assert rows.shape == (20000, 5)
    x = np.abs(rows[:, 0].astype(np.uint32).view(np.float32))
    h = rows[:, 1].astype(np.uint16)
    finite = ~np.isnan(x)
    with np.errstate(over="ignore"):
        near = x[finite].astype(np.float16)  # round to nearest even ...
        low = near.astype(np.float32) < x[finite]  # ... then one step up where that fell below x
        want = np.where(low, np.nextafter(near, np.float16(np.inf)), near)
    assert np.array_equal(h[finite], want.view(np.uint16))
    assert np.all(h[~finite] == 0x7c00)  # NaN extent: unbounded side
    dec = h.view(np.float16).astype(np.float32).view(np.uint32)
    for col in (2, 3, 4):
        assert np.array_equal(rows[:, col].astype(np.uint32), dec)
    assert np.all(dec.view(np.float32)[finite] >= x[finite])
