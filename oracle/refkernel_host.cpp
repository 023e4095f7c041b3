// oracle/refkernel_host.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Headless harness for the reference's OWN CUDA kernel: Kernel.cu is compiled UNMODIFIED for sm_100a
// (oracle/Makefile, release flags -O3 -use_fast_math of CudaRayTracer/CMakeLists.txt:36) and called
// through its own extern "C" launchers exactly as CudaLayer does (CudaLayer.cpp:93-101,372-375).
// This file plays the part of CudaLayer.cpp: compiled by the HOST compiler with the reference headers,
// it builds the scene in cudaMallocManaged memory through the reference's classes and times LaunchKernel.
// It is the "reference kernel on one B200" comparator of BASELINE.md §3 (R) and runs only on the GPU box.

#include <cuda_runtime.h>
#include <curand_kernel.h>
#include <thrust/device_vector.h>
#include <thrust/partition.h>
#include <thrust/remove.h>
#include <thrust/sort.h>

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unordered_map>
#include <vector>

#include "Hittables/Hittable.cuh"
#include "Hittables/Material.cuh"
#include "Utils/SharedStructs.h"

#include "../include/crt.h"

std::shared_ptr<spdlog::logger> Log::s_Logger;
// Core/Log.cpp wires an ImGui console sink; headless, a sink-less logger is enough (only reached on error paths).
void Log::Init()
{
    if (!s_Logger) s_Logger = std::make_shared<spdlog::logger>("APP");
}

// The reference's launchers (Kernel.cu:178-204), declared as CudaLayer.cpp:12-19 declares them.
extern "C" void LaunchKernel(unsigned int* pos, unsigned int image_width, unsigned int image_height,
                             const unsigned int samples_per_pixel, const unsigned int max_depth, Hittable* world,
                             curandState* d_rand_state, InputStruct inputs);
extern "C" void LaunchRandInit(curandState* d_rand_state2);
extern "C" void LaunchRenderInit(dim3 grid, dim3 block, unsigned int window_width, unsigned int window_height,
                                 curandState* d_rand_state);
// oracle/refkernel_ids.cu
extern "C" int refkernel_ids_launch(Hittable* world, const char* slab, size_t list_bytes, size_t record_bytes, size_t n,
                                    InputStruct inputs, unsigned width, unsigned height, int* d_out);
// the same translation unit compiled with IEEE arithmetic (-fmad=false, no -use_fast_math)
extern "C" int refkernel_ids_launch_ieee(Hittable* world, const char* slab, size_t list_bytes, size_t record_bytes, size_t n,
                                         InputStruct inputs, unsigned width, unsigned height, int* d_out);

namespace {

static void* managed_zeroed(size_t n)
{
    void* p = nullptr;
    if (cudaMallocManaged(&p, n ? n : 1) != cudaSuccess) return nullptr;
    memset(p, 0, n);
    return p;
}
#define REF_ALLOC(n) managed_zeroed(n)
#define REF_FREE(p) cudaFree(p)
#define REF_COPY_IMAGES 1
#include "ref_scene.inc"

} // namespace

extern "C" {

// Renders `reps` frames with the reference kernel; ms[i] = host wall-clock of the i-th LaunchKernel call (it ends
// with cudaDeviceSynchronize, Kernel.cu:190). One untimed warm-up launch pages the managed slab onto the GPU.
// out_rgba8 (host, W*H) receives the last frame; pixels outside the truncated grid (Kernel.cu:183-184) stay 0.
int refkernel_render(const crt_scene_desc* scene, const crt_input_struct* cam, uint32_t width, uint32_t height, uint32_t spp,
                     uint32_t max_depth, int reps, uint32_t* out_rgba8, float* ms)
{
    Log::Init();
    RefScene rs;
    if (!BuildScene(scene, rs)) { FreeScene(rs); return 1; }
    InputStruct in;
    static_assert(sizeof(InputStruct) == sizeof(crt_input_struct), "InputStruct layout");
    memcpy(&in, cam, sizeof(in));
    unsigned int* d_pos = nullptr;
    curandState* d_rand = nullptr;
    const size_t npx = (size_t)width * height;
    int rc = 0;
    if (cudaMalloc(&d_pos, npx * 4) != cudaSuccess || cudaMalloc(&d_rand, npx * sizeof(curandState)) != cudaSuccess) rc = 2;
    if (!rc) {
        cudaMemset(d_pos, 0, npx * 4);
        dim3 block(16, 16, 1), grid(width / block.x, height / block.y, 1); // CudaLayer.cpp:97-100
        LaunchRenderInit(grid, block, width, height, d_rand);
        LaunchKernel(d_pos, width, height, 1, max_depth, rs.world, d_rand, in); // warm-up
        for (int i = 0; i < reps; i++) {
            auto t0 = std::chrono::steady_clock::now();
            LaunchKernel(d_pos, width, height, spp, max_depth, rs.world, d_rand, in);
            auto t1 = std::chrono::steady_clock::now();
            if (ms) ms[i] = std::chrono::duration<float, std::milli>(t1 - t0).count();
        }
        if (cudaGetLastError() != cudaSuccess) rc = 3;
        if (out_rgba8 && cudaMemcpy(out_rgba8, d_pos, npx * 4, cudaMemcpyDeviceToHost) != cudaSuccess) rc = 4;
    }
    cudaFree(d_pos);
    cudaFree(d_rand);
    FreeScene(rs);
    return rc;
}

// Primary-hit hittable indices (xi = 0.5) computed by the reference's own device BVHNode::Hit; ieee = 0: the reference's release
// flags (fast math), 1: IEEE arithmetic.
static int primary_ids(const crt_scene_desc* scene, const crt_input_struct* cam, uint32_t width, uint32_t height, int32_t* out_ids, int ieee);
int refkernel_primary_ids(const crt_scene_desc* scene, const crt_input_struct* cam, uint32_t width, uint32_t height, int32_t* out_ids)
{
    return primary_ids(scene, cam, width, height, out_ids, 0);
}
int refkernel_primary_ids_ieee(const crt_scene_desc* scene, const crt_input_struct* cam, uint32_t width, uint32_t height, int32_t* out_ids)
{
    return primary_ids(scene, cam, width, height, out_ids, 1);
}
static int primary_ids(const crt_scene_desc* scene, const crt_input_struct* cam, uint32_t width, uint32_t height, int32_t* out_ids, int ieee)
{
    Log::Init();
    RefScene rs;
    if (!BuildScene(scene, rs)) { FreeScene(rs); return 1; }
    InputStruct in;
    memcpy(&in, cam, sizeof(in));
    int* d_out = nullptr;
    const size_t npx = (size_t)width * height;
    int rc = 0;
    if (cudaMalloc(&d_out, npx * 4) != cudaSuccess) rc = 2;
    if (!rc) rc = (ieee ? refkernel_ids_launch_ieee : refkernel_ids_launch)(rs.world, rs.slab, rs.list_bytes, rs.record_bytes, rs.n, in, width, height, d_out);
    if (!rc && cudaMemcpy(out_ids, d_out, npx * 4, cudaMemcpyDeviceToHost) != cudaSuccess) rc = 4;
    cudaFree(d_out);
    FreeScene(rs);
    return rc;
}

} // extern "C"
