// oracle/ref_abi_probe.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// Compile-time check of cudaraytracer_b200/csrc/ref_abi.h (the layout mirror libcrt.so walks the caller's scene graph with)
// against the reference's OWN classes: every sizeof and every offsetof the flattener relies on. Compiles to an empty object;
// a layout drift in either tree is a compile error. Built by `make -C oracle ref` and by tests/test_ref_abi.py.
#include <cstddef>

#include "Hittables/Hittable.cuh"
#include "Hittables/Material.cuh"
#include "Hittables/Texture.cuh"
#include "Hittables/AABB.cuh"

#include "../cudaraytracer_b200/csrc/ref_abi.h"

#define SAME_SIZE(A, B) static_assert(sizeof(A) == sizeof(B), "sizeof " #A " != sizeof " #B)
#define SAME_OFF(A, fa, B, fb) static_assert(offsetof(A, fa) == offsetof(B, fb), "offsetof " #A "::" #fa " != " #B "::" #fb)

SAME_SIZE(::Vec3, ref_abi::Vec3);
SAME_SIZE(::AABB, ref_abi::AABB);
SAME_OFF(::AABB, minimum, ref_abi::AABB, minimum);
SAME_OFF(::AABB, maximum, ref_abi::AABB, maximum);

SAME_SIZE(::Hittable, ref_abi::Hittable);
SAME_OFF(::Hittable, type, ref_abi::Hittable, type);
SAME_OFF(::Hittable, isActive, ref_abi::Hittable, isActive);
SAME_OFF(::Hittable, Object, ref_abi::Hittable, Object);
static_assert(sizeof(::Hittable::ObjectUnion) == sizeof(void*), "ObjectUnion is one pointer");

SAME_SIZE(::Sphere, ref_abi::Sphere);
SAME_OFF(::Sphere, center, ref_abi::Sphere, center);
SAME_OFF(::Sphere, radius, ref_abi::Sphere, radius);
SAME_OFF(::Sphere, mat_ptr, ref_abi::Sphere, mat_ptr);

SAME_SIZE(::XYRect, ref_abi::Rect);
SAME_SIZE(::XZRect, ref_abi::Rect);
SAME_SIZE(::YZRect, ref_abi::Rect);
SAME_OFF(::XYRect, center, ref_abi::Rect, center);
SAME_OFF(::XYRect, width, ref_abi::Rect, width);
SAME_OFF(::XYRect, height, ref_abi::Rect, height);
SAME_OFF(::XYRect, mat_ptr, ref_abi::Rect, mat_ptr);
SAME_OFF(::XZRect, center, ref_abi::Rect, center);
SAME_OFF(::XZRect, width, ref_abi::Rect, width);
SAME_OFF(::XZRect, height, ref_abi::Rect, height);
SAME_OFF(::XZRect, mat_ptr, ref_abi::Rect, mat_ptr);
SAME_OFF(::YZRect, center, ref_abi::Rect, center);
SAME_OFF(::YZRect, width, ref_abi::Rect, width);
SAME_OFF(::YZRect, height, ref_abi::Rect, height);
SAME_OFF(::YZRect, mat_ptr, ref_abi::Rect, mat_ptr);

SAME_SIZE(::BVHNode, ref_abi::BVHNode);
SAME_OFF(::BVHNode, box, ref_abi::BVHNode, box);
SAME_OFF(::BVHNode, left, ref_abi::BVHNode, left);
SAME_OFF(::BVHNode, right, ref_abi::BVHNode, right);

SAME_SIZE(::HittableList, ref_abi::HittableList);
SAME_OFF(::HittableList, list, ref_abi::HittableList, list);
SAME_OFF(::HittableList, list_size, ref_abi::HittableList, list_size);

SAME_SIZE(::Material, ref_abi::Material);
SAME_OFF(::Material, type, ref_abi::Material, type);
SAME_OFF(::Material, Object, ref_abi::Material, Object);
SAME_SIZE(::Lambertian, ref_abi::Lambertian);
SAME_OFF(::Lambertian, albedo, ref_abi::Lambertian, albedo);
SAME_SIZE(::Metal, ref_abi::Metal);
SAME_OFF(::Metal, albedo, ref_abi::Metal, albedo);
SAME_OFF(::Metal, fuzz, ref_abi::Metal, fuzz);
SAME_SIZE(::Dielectric, ref_abi::Dielectric);
SAME_OFF(::Dielectric, ir, ref_abi::Dielectric, ir);
SAME_SIZE(::DiffuseLight, ref_abi::DiffuseLight);
SAME_OFF(::DiffuseLight, albedo, ref_abi::DiffuseLight, albedo);
SAME_OFF(::DiffuseLight, light_intensity, ref_abi::DiffuseLight, light_intensity);

SAME_SIZE(::Texture, ref_abi::Texture);
SAME_OFF(::Texture, type, ref_abi::Texture, type);
SAME_OFF(::Texture, Object, ref_abi::Texture, Object);
SAME_SIZE(::Constant, ref_abi::Constant);
SAME_OFF(::Constant, color, ref_abi::Constant, color);
SAME_SIZE(::Checker, ref_abi::Checker);
SAME_OFF(::Checker, odd, ref_abi::Checker, odd);
SAME_OFF(::Checker, even, ref_abi::Checker, even);
SAME_SIZE(::Image, ref_abi::Image);
SAME_OFF(::Image, data, ref_abi::Image, data);
SAME_OFF(::Image, width, ref_abi::Image, width);
SAME_OFF(::Image, height, ref_abi::Image, height);

static_assert((int)HittableType::SPHERE == ref_abi::HT_SPHERE && (int)HittableType::XYRECT == ref_abi::HT_XYRECT &&
              (int)HittableType::XZRECT == ref_abi::HT_XZRECT && (int)HittableType::YZRECT == ref_abi::HT_YZRECT &&
              (int)HittableType::HITTABLELIST == ref_abi::HT_HITTABLELIST && (int)HittableType::BVHNODE == ref_abi::HT_BVHNODE,
              "HittableType values");
