// oracle/refkernel_ids.cu — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// A tiny kernel of OURS that calls the REFERENCE'S device BVHNode::Hit (Hittable.cuh:387-439) for one
// xi = 0.5 primary ray per pixel and maps rec.mat_ptr back to the hittable-list index (each record of the
// slab owns its Material, so the pointer identifies the hittable). Ray generation restates Kernel.cu:130-146.
// Compiled twice (oracle/Makefile): with the reference's release flags (-use_fast_math: the GPU flavour the reference ships,
// entry refkernel_ids_launch) and with IEEE arithmetic (-fmad=false, no fast math: entry refkernel_ids_launch_ieee), the
// flavour whose IDs the new path must reproduce BIT FOR BIT (SURVEY.md §7 hard parts (iii), north_star).
#ifndef REFIDS_KERNEL
#define REFIDS_KERNEL RefPrimaryIds
#define REFIDS_LAUNCH refkernel_ids_launch
#endif
#include <cuda_runtime.h>
#include <device_launch_parameters.h>
#include <float.h>

#include "Hittables/Hittable.cuh"
#include "Hittables/Material.cuh"
#include "Utils/SharedStructs.h"

__global__ void REFIDS_KERNEL(Hittable* world, const char* slab, size_t list_bytes, size_t record_bytes, size_t n, InputStruct inputs,
                              unsigned width, unsigned height, int* out)
{
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= width || y >= height) return;
    Vec3 origin = Vec3(inputs.origin_x, inputs.origin_y, inputs.origin_z);
    Vec3 forwardV = Vec3(inputs.orientation_x, inputs.orientation_y, inputs.orientation_z);
    Vec3 upV = Vec3(inputs.up_x, inputs.up_y, inputs.up_z);
    Vec3 rightV = Normalize(Cross(upV, forwardV));
    Vec3 center = Vec3(width / 2.0f, height / 2.0f, 0.0f);
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
        size_t off = (const char*)rec.mat_ptr - (slab + list_bytes);
        id = (int)(off / record_bytes);
        if ((size_t)id >= n) id = -2;
    }
    out[y * width + x] = id;
}

extern "C" int REFIDS_LAUNCH(Hittable* world, const char* slab, size_t list_bytes, size_t record_bytes, size_t n,
                                    InputStruct inputs, unsigned width, unsigned height, int* d_out)
{
    dim3 block(16, 16, 1), grid((width + 15) / 16, (height + 15) / 16, 1);
    REFIDS_KERNEL<<<grid, block>>>(world, slab, list_bytes, record_bytes, n, inputs, width, height, d_out);
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 3;
}
