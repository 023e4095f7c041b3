"""The layout mirror libcrt.so walks the caller's scene graph with (cudaraytracer_b200/csrc/ref_abi.h) against the reference's own
classes: oracle/ref_abi_probe.cpp includes both and static_asserts every sizeof / offsetof / enum value the flattener relies on
(Hittable.cuh:47-78,128-137,296-301,532-538, Material.cuh, Texture.cuh, AABB.cuh). Compiling it IS the test. Needs the
reference tree, i.e. this container; on the GPU box the prebuilt oracle/_ref/ref_abi_probe.o is the record of the check."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"


@pytest.mark.skipif(not os.path.exists(os.path.join(REF, "CudaRayTracer", "src", "Hittables", "Hittable.cuh")), reason="needs /root/reference")
def test_ref_abi_mirror_matches_the_reference_headers(tmp_path):
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = ["g++", "-std=c++17", "-O0", "-DRT_RELEASE", "-DTHRUST_DEVICE_SYSTEM=THRUST_DEVICE_SYSTEM_CPP", "-w", "-Wno-invalid-offsetof",
           "-I", os.path.join(cuda, "include"), "-I", os.path.join(REF, "CudaRayTracer", "src"), "-I", os.path.join(REF, "vendor", "spdlog", "include"),
           "-c", os.path.join(ROOT, "oracle", "ref_abi_probe.cpp"), "-o", str(tmp_path / "probe.o")]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=os.path.join(ROOT, "oracle"))
    assert r.returncode == 0, r.stderr[-4000:]


def test_the_probe_really_checks(tmp_path):
    """A mirror with a shifted field must NOT compile (guards against the probe silently asserting nothing)."""
    if not os.path.exists(os.path.join(REF, "CudaRayTracer", "src", "Hittables", "Hittable.cuh")):
        pytest.skip("needs /root/reference")
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    bad_dir = tmp_path / "cudaraytracer_b200" / "csrc"
    bad_dir.mkdir(parents=True)
    src = open(os.path.join(ROOT, "cudaraytracer_b200", "csrc", "ref_abi.h")).read()
    (bad_dir / "ref_abi.h").write_text(src.replace("    Vec3 center;\n    float radius;\n    Material* mat_ptr;", "    Vec3 center;\n    Material* mat_ptr;\n    float radius;"))
    (tmp_path / "oracle").mkdir()
    (tmp_path / "oracle" / "ref_abi_probe.cpp").write_text(open(os.path.join(ROOT, "oracle", "ref_abi_probe.cpp")).read())
    cmd = ["g++", "-std=c++17", "-O0", "-DRT_RELEASE", "-DTHRUST_DEVICE_SYSTEM=THRUST_DEVICE_SYSTEM_CPP", "-w", "-Wno-invalid-offsetof",
           "-I", os.path.join(cuda, "include"), "-I", os.path.join(REF, "CudaRayTracer", "src"), "-I", os.path.join(REF, "vendor", "spdlog", "include"),
           "-c", str(tmp_path / "oracle" / "ref_abi_probe.cpp"), "-o", str(tmp_path / "probe.o")]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=str(tmp_path / "oracle"))
    assert r.returncode != 0 and "static assertion" in r.stderr
