// ref_abi.h — layout mirror of the reference's scene-graph records, so that libcrt.so can walk the
// pointer graph a CudaRayTracer host hands to LaunchKernel (Kernel.cu:178-181) without including the
// reference's headers. Plain PODs with the same size, alignment and field offsets as the classes in
//   Hittables/Hittable.cuh:47-67 (Hittable), :69-78 (Sphere), :128-137 (XYRect; XZRect/YZRect identical),
//   :296-301,467 (BVHNode), Hittables/AABB.cuh:5-9, Hittables/Material.cuh:19-32,34-41,65-73,97-104,148-156,
//   Hittables/Texture.cuh:18-30,32-40,47-56,70-81,106-108, Utils/Math.cuh:94 (Vec3 = float e[3]).
// oracle/ref_abi_probe.cpp includes BOTH this file and the reference headers and static_asserts every sizeof / offsetof
// below against the real classes; tests/test_ref_abi.py compiles it (where /root/reference is present), and so does
// `make -C oracle ref`.
#pragma once

#include <cstddef>
#include <cstdint>

namespace ref_abi {

struct Vec3 { float e[3]; };

struct AABB { Vec3 minimum, maximum; };

struct Hittable
{
    int32_t type;   // HittableType
    bool isActive;
    void** Object;  // -> ObjectUnion (one pointer) -> Sphere / XYRect / ... / BVHNode
};

struct Material;

struct Sphere
{
    Vec3 center;
    float radius;
    Material* mat_ptr;
};

struct Rect // XYRect, XZRect, YZRect share this layout
{
    Vec3 center;
    float width;
    float height;
    Material* mat_ptr;
};

struct BVHNode
{
    AABB box;
    Hittable* left;
    Hittable* right;
    char* memory;
};

struct HittableList // Hittable.cuh:532-581: every element is treated as a BVHNode hittable (its Hit calls list[i]->Object->bvh_node)
{
    Hittable** list;
    int list_size;
};

struct Texture
{
    int32_t type;  // TextureType
    void** Object; // -> ObjectUnion -> Constant / Checker / Image
};

struct Constant
{
    Vec3 color;
    float padding;
};

struct Checker
{
    Constant* odd;
    Constant* even;
};

struct Image
{
    unsigned char* data;
    const char* path;
    int width, height;
    int bytes_per_scanline;
};

struct Material
{
    int32_t type;  // MaterialType
    void** Object; // -> ObjectUnion -> Lambertian / Metal / Dielectric / DiffuseLight
};

struct Lambertian { Texture* albedo; };
struct Metal { Texture* albedo; float fuzz; };
struct Dielectric { float ir; };
struct DiffuseLight { Texture* albedo; int light_intensity; };

enum { HT_SPHERE = 0, HT_XYRECT = 1, HT_XZRECT = 2, HT_YZRECT = 3, HT_HITTABLELIST = 4, HT_BVHNODE = 5 };

static_assert(sizeof(Hittable) == 16 && offsetof(Hittable, Object) == 8, "Hittable");
static_assert(sizeof(Sphere) == 24 && offsetof(Sphere, mat_ptr) == 16, "Sphere");
static_assert(sizeof(Rect) == 32 && offsetof(Rect, mat_ptr) == 24, "Rect");
static_assert(sizeof(BVHNode) == 48 && offsetof(BVHNode, left) == 24, "BVHNode");
static_assert(sizeof(HittableList) == 16 && offsetof(HittableList, list_size) == 8, "HittableList");
static_assert(sizeof(Material) == 16 && sizeof(Texture) == 16, "Material/Texture");
static_assert(sizeof(Constant) == 16 && sizeof(Checker) == 16 && sizeof(Image) == 32, "textures");
static_assert(sizeof(Metal) == 16 && sizeof(DiffuseLight) == 16, "materials");

} // namespace ref_abi
