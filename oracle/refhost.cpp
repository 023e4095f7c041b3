// oracle/refhost.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// The reference's own hot-path headers (Hittable.cuh, Material.cuh, Texture.cuh, AABB.cuh, Math.cuh,
// Ray.cuh, SharedStructs.h) compiled FOR THE HOST with g++, straight from /root/reference (nothing is
// copied into this repo). The host compiler ignores __device__, and cuRAND's XORWOW device API has a
// host code path, so Sphere::Hit, *Rect::Hit, AABB::Hit, the BVHNode constructor and BVHNode::Hit, every
// Scatter/Emitted, every Texture::value, Reflect/Refract and curand_uniform below are the REFERENCE'S
// code. Only what lives in Kernel.cu and depends on threadIdx is restated here, operation for operation:
//   color()            Kernel.cu:30-80
//   Kernel per-pixel   Kernel.cu:113-157
//   RgbToInt           Kernel.cu:12-19
//   RenderInit seeding Kernel.cu:175
// Scene records are laid out the way CudaLayer::GenerateWorld does it (CudaLayer.cpp:108-129,177-195):
// one slab, each hittable record sized for the largest material/texture variant.
//
// Built by oracle/Makefile into oracle/_ref/libcrt_refhost.so (git-ignored, travels to the GPU box).
// Used by tests/ to pin oracle/oracle.c, and by bench.py as cpu_baseline kind "reference".

#include <cuda_runtime.h>
#include <curand_kernel.h>
#include <thrust/device_vector.h>
#include <thrust/partition.h>
#include <thrust/remove.h>
#include <thrust/sort.h>

#include <atomic>
#include <cfloat>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <unordered_map>
#include <vector>

// The BVHNode constructor allocates with cudaMallocManaged (Hittable.cuh:348); there is no GPU here.
static inline cudaError_t refhost_calloc(void** p, size_t n)
{
    *p = calloc(1, n);
    return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
template <class T> static inline cudaError_t refhost_calloc_t(T** p, size_t n)
{
    return refhost_calloc((void**)p, n);
}
#define cudaMallocManaged(p, n) refhost_calloc_t((p), (n))
#define cudaFree(p) (free(p), cudaSuccess)

#include "Hittables/Hittable.cuh"
#include "Hittables/Material.cuh"
#include "Utils/SharedStructs.h"

#undef cudaMallocManaged
#undef cudaFree

#include "../include/crt.h"

std::shared_ptr<spdlog::logger> Log::s_Logger;

namespace {

#define REF_ALLOC(n) calloc(1, (n))
#define REF_FREE(p) free(p)
#include "ref_scene.inc"

InputStruct ToInputs(const crt_input_struct* c)
{
    InputStruct in;
    static_assert(sizeof(InputStruct) == sizeof(crt_input_struct), "InputStruct layout");
    memcpy(&in, c, sizeof(in));
    return in;
}

// Kernel.cu:12-19
inline int RgbToIntHost(float r, float g, float b)
{
    r = Clamp(r, 0.0f, 255.0f);
    g = Clamp(g, 0.0f, 255.0f);
    b = Clamp(b, 0.0f, 255.0f);
    float a = 255.0f;
    return (int(a) << 24) | (int(b) << 16) | (int(g) << 8) | int(r);
}

// Kernel.cu:30-80, plus a ray counter (one per closest-hit query) that does not touch the arithmetic.
inline Vec3 ColorHost(const Ray& r, Hittable* world, int max_depth, curandState* local_rand_state, const InputStruct& inputs,
                      uint64_t& rays)
{
    Ray cur_ray = r;
    Vec3 cur_attenuation = Vec3(1.0f, 1.0f, 1.0f);
    Vec3 black_color = Vec3(0.0f, 0.0f, 0.0f);
    HitRecord rec;

    for (int i = 0; i < max_depth; i++) {
        rays++;
        if (!world->Object->bvh_node->Hit(cur_ray, 0.001f, FLT_MAX, rec)) {
            Vec3 unit_direction = UnitVector(cur_ray.Direction());
            float t = 0.5f * (unit_direction.y() + 1.0f);
            Vec3 c = (1.0f - t) * inputs.background_start + t * inputs.background_end;
            return cur_attenuation * c;
        }
        else {
            Vec3 emitted = Vec3(0.0f, 0.0f, 0.0f);
            Ray scattered;
            Vec3 attenuation;
            switch (rec.mat_ptr->type) {
            case MaterialType::LAMBERTIAN:
                if (!rec.mat_ptr->Object->lambertian->Scatter(cur_ray, rec, attenuation, scattered, local_rand_state))
                    return emitted * cur_attenuation;
                break;
            case MaterialType::METAL:
                if (!rec.mat_ptr->Object->metal->Scatter(cur_ray, rec, attenuation, scattered, local_rand_state))
                    return emitted * cur_attenuation;
                break;
            case MaterialType::DIELECTRIC:
                if (!rec.mat_ptr->Object->dielectric->Scatter(cur_ray, rec, attenuation, scattered, local_rand_state))
                    return emitted * cur_attenuation;
                break;
            case MaterialType::DIFFUSELIGHT:
                emitted = rec.mat_ptr->Object->diffuse_light->Emitted(rec.u, rec.v, rec.p);
                return emitted * cur_attenuation;
            default:
                return black_color;
            }
            cur_attenuation = attenuation * cur_attenuation;
            cur_ray = scattered;
        }
    }
    return black_color;
}

struct Job
{
    RefScene* rs;
    InputStruct inputs;
    unsigned width, height, spp, max_depth;
    int jitter_center;
    uint32_t* out_rgba8;
    float* out_mean;
    int32_t* out_ids;
    std::atomic<unsigned> next_row{0};
    std::atomic<uint64_t> rays{0};
};

// Kernel.cu:113-157 for one pixel.
void PixelHost(Job& J, int x, int y, uint64_t& rays)
{
    const unsigned width = J.width, height = J.height;
    const InputStruct& inputs = J.inputs;
    Hittable* world = J.rs->world;
    unsigned int pixel_index = (y * width + x);

    curandState local_rand_state;
    curand_init(1984 + pixel_index, 0, 0, &local_rand_state); // Kernel.cu:175

    Vec3 col = Vec3(0.0f, 0.0f, 0.0f);
    Vec3 origin = Vec3(inputs.origin_x, inputs.origin_y, inputs.origin_z);
    Vec3 forwardV = Vec3(inputs.orientation_x, inputs.orientation_y, inputs.orientation_z);
    Vec3 upV = Vec3(inputs.up_x, inputs.up_y, inputs.up_z);
    Vec3 rightV = Normalize(Cross(upV, forwardV));
    Vec3 center = Vec3(width / 2.0f, height / 2.0f, 0.0f);

    if (J.out_ids) {
        float u = (float)((x - center.x()) + 0.5f) / (float)(width);
        float v = (float)((center.y() - y) + 0.5f) / (float)(width);
        Vec3 distFromCenter = (u * rightV) + (v * upV);
        Vec3 startPos = (inputs.near_plane * distFromCenter) + origin + (inputs.fov * forwardV);
        Vec3 secondPlanePos = (inputs.far_plane * distFromCenter) + ((1.0f / inputs.fov * 10.0f) * forwardV) + origin;
        Vec3 dirVector = Normalize(secondPlanePos - startPos);
        Ray r = Ray(startPos, dirVector);
        HitRecord rec;
        int id = -1;
        if (world->Object->bvh_node->Hit(r, 0.001f, FLT_MAX, rec)) {
            auto it = J.rs->mat_to_index.find(rec.mat_ptr);
            id = it == J.rs->mat_to_index.end() ? -2 : it->second;
        }
        J.out_ids[pixel_index] = id;
    }
    if (!J.out_rgba8 && !J.out_mean)
        return;

    for (unsigned s = 0; s < J.spp; s++) {
        float j1 = J.jitter_center ? 0.5f : curand_uniform(&local_rand_state);
        float j2 = J.jitter_center ? 0.5f : curand_uniform(&local_rand_state);
        float u = (float)((x - center.x()) + j1) / (float)(width);
        float v = (float)((center.y() - y) + j2) / (float)(width);
        Vec3 distFromCenter = (u * rightV) + (v * upV);
        Vec3 startPos = (inputs.near_plane * distFromCenter) + origin + (inputs.fov * forwardV);
        Vec3 secondPlanePos = (inputs.far_plane * distFromCenter) + ((1.0f / inputs.fov * 10.0f) * forwardV) + origin;
        Vec3 dirVector = Normalize(secondPlanePos - startPos);
        Ray r = Ray(startPos, dirVector);
        col = col + ColorHost(r, world, J.max_depth, &local_rand_state, inputs, rays);
    }
    col = col / (float)(J.spp);
    if (J.out_mean) {
        J.out_mean[3 * (size_t)pixel_index + 0] = col.x();
        J.out_mean[3 * (size_t)pixel_index + 1] = col.y();
        J.out_mean[3 * (size_t)pixel_index + 2] = col.z();
    }
    col.e[0] = 255.0f * sqrtf(col.x());
    col.e[1] = 255.0f * sqrtf(col.y());
    col.e[2] = 255.0f * sqrtf(col.z());
    if (J.out_rgba8)
        J.out_rgba8[pixel_index] = (uint32_t)RgbToIntHost(col.x(), col.y(), col.z());
}

void Worker(Job* J)
{
    uint64_t rays = 0;
    for (;;) {
        unsigned y = J->next_row.fetch_add(1);
        if (y >= J->height)
            break;
        for (unsigned x = 0; x < J->width; x++)
            PixelHost(*J, (int)x, (int)y, rays);
    }
    J->rays.fetch_add(rays);
}

} // namespace

extern "C" {

// Renders every pixel (the W/16 x H/16 truncation is a property of LaunchKernel, Kernel.cu:183-184, not of the math).
// out_rgba8: W*H packed pixels or NULL. out_mean: W*H*3 floats (col / spp, before gamma) or NULL.
// out_ids: W*H hittable-list indices of the xi=0.5 primary ray (-1 miss) or NULL. Returns 0 on success.
int refhost_render(const crt_scene_desc* scene, const crt_input_struct* cam, uint32_t width, uint32_t height, uint32_t spp,
                   uint32_t max_depth, int jitter_center, int n_threads, uint32_t* out_rgba8, float* out_mean,
                   int32_t* out_ids, uint64_t* out_rays)
{
    RefScene rs;
    if (!BuildScene(scene, rs)) {
        FreeScene(rs);
        return 1;
    }
    Job J;
    J.rs = &rs;
    J.inputs = ToInputs(cam);
    J.width = width;
    J.height = height;
    J.spp = spp;
    J.max_depth = max_depth;
    J.jitter_center = jitter_center;
    J.out_rgba8 = out_rgba8;
    J.out_mean = out_mean;
    J.out_ids = out_ids;
    if (n_threads < 1)
        n_threads = (int)std::thread::hardware_concurrency();
    if (n_threads < 1)
        n_threads = 1;
    std::vector<std::thread> th;
    for (int i = 1; i < n_threads; i++)
        th.emplace_back(Worker, &J);
    Worker(&J);
    for (auto& t : th)
        t.join();
    if (out_rays)
        *out_rays = J.rays.load();
    FreeScene(rs);
    return 0;
}

// Reference scene graph in HOST memory for driving the drop-in LaunchKernel of libcrt.so from tests
// (the graph is what CudaLayer hands over; cudaMallocManaged memory is host-readable in the same way).
void* refhost_world_create(const crt_scene_desc* scene)
{
    RefScene* rs = new RefScene();
    if (!BuildScene(scene, *rs)) {
        FreeScene(*rs);
        delete rs;
        return nullptr;
    }
    return rs;
}
void* refhost_world_root(void* handle)
{
    return handle ? (void*)((RefScene*)handle)->world : nullptr;
}
void* refhost_world_list(void* handle)
{
    return handle ? (void*)((RefScene*)handle)->list : nullptr;
}
// The same world behind a HittableList root (Hittable.cuh:532-581): a list whose single element is the BVH root hittable.
// For the drop-in test of crt_scene_from_reference / LaunchKernel with a list root. Freed with the process.
void* refhost_world_list_root(void* handle)
{
    if (!handle) return nullptr;
    RefScene* rs = (RefScene*)handle;
    struct Wrap { char h[sizeof(Hittable)]; Hittable::ObjectUnion u; char l[sizeof(HittableList)]; Hittable* elems[1]; };
    Wrap* w = (Wrap*)calloc(1, sizeof(Wrap));
    if (!w) return nullptr;
    w->elems[0] = rs->world;
    HittableList* l = new (w->l) HittableList(w->elems, 1);
    w->u.hittable_list = l;
    Hittable* h = (Hittable*)w->h;
    h->type = HittableType::HITTABLELIST;
    h->isActive = true;
    h->Object = &w->u;
    return h;
}
void refhost_world_destroy(void* handle)
{
    if (!handle) return;
    FreeScene(*(RefScene*)handle);
    delete (RefScene*)handle;
}

// XORWOW known-answer hook: n uniforms from curand_init(seed, 0, 0).
void refhost_curand_uniform(uint64_t seed, int n, float* out)
{
    curandState s;
    curand_init(seed, 0, 0, &s);
    for (int i = 0; i < n; i++)
        out[i] = curand_uniform(&s);
}

// Struct sizes of the reference ABI as this compiler sees them.
void refhost_struct_sizes(int32_t* out)
{
    int i = 0;
    out[i++] = sizeof(Hittable);
    out[i++] = sizeof(Sphere);
    out[i++] = sizeof(BVHNode);
    out[i++] = sizeof(Material);
    out[i++] = sizeof(Lambertian);
    out[i++] = sizeof(Metal);
    out[i++] = sizeof(Dielectric);
    out[i++] = sizeof(DiffuseLight);
    out[i++] = sizeof(Texture);
    out[i++] = sizeof(Constant);
    out[i++] = sizeof(Checker);
    out[i++] = sizeof(Image);
    out[i++] = sizeof(HitRecord);
    out[i++] = sizeof(curandState);
    out[i++] = sizeof(XYRect);
    out[i++] = sizeof(InputStruct);
}

// Single-primitive probes used to pin the restatement function by function.
// Sphere::Hit (Hittable.cuh:80-110): returns hit flag; out = t, px,py,pz, nx,ny,nz, u, v.
int refhost_sphere_hit(const float center[3], float radius, const float ro[3], const float rd[3], float t_min, float t_max,
                       float* out)
{
    Sphere s(Vec3(center[0], center[1], center[2]), radius, nullptr);
    Ray r(Vec3(ro[0], ro[1], ro[2]), Vec3(rd[0], rd[1], rd[2]));
    HitRecord rec;
    bool h = s.Hit(r, t_min, t_max, rec);
    if (h) {
        out[0] = rec.t;
        out[1] = rec.p.x(); out[2] = rec.p.y(); out[3] = rec.p.z();
        out[4] = rec.normal.x(); out[5] = rec.normal.y(); out[6] = rec.normal.z();
        out[7] = rec.u; out[8] = rec.v;
    }
    return h ? 1 : 0;
}

// AABB::Hit (AABB.cuh:30-50)
int refhost_aabb_hit(const float lo[3], const float hi[3], const float ro[3], const float rd[3], float t_min, float t_max)
{
    AABB b(Vec3(lo[0], lo[1], lo[2]), Vec3(hi[0], hi[1], hi[2]));
    Ray r(Vec3(ro[0], ro[1], ro[2]), Vec3(rd[0], rd[1], rd[2]));
    return b.Hit(r, t_min, t_max) ? 1 : 0;
}

} // extern "C"
